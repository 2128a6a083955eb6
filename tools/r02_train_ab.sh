#!/bin/bash
# same-box A/B of the training step over library builds: LIBS="name ..." ("tree" = working tree)
o=gpurun_out
for n in $LIBS; do
  lib=build/libt2b200_$n.so; [ $n = tree ] && lib=tacotron2_b200/libt2b200.so
  echo "== $n"
  T2B200_LIB=$lib python tools/train_probe.py 4 2>&1 | tail -2 | head -1 | cut -c1-330
done 2>&1 | tee $o/${TAG:-r02_train}_ab.log
