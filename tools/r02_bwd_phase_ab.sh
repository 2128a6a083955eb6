#!/bin/bash
# same-box A/B of decoder_bwd_kernel phase times over library builds: LIBS="name ..." ("tree" = working tree)
o=gpurun_out
for n in $LIBS; do
  lib=build/libt2b200_$n.so; [ $n = tree ] && lib=tacotron2_b200/libt2b200.so
  echo "== $n"
  T2B200_LIB=$lib python bench.py --train-only --steps 3 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); t=d.get('train',d)
print(round(t['ms_per_step'],2), round(t['roofline']['kernel_ms'],2), round(t['roofline']['frac'],4), round(t['decoder_fwd_kernel_ms'],2)); print(t['decoder_bwd_phase_us_per_frame'])"
done 2>&1 | tee $o/${TAG:-r02_bwd_phase}_ab.log
