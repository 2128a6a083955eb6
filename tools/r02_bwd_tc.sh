#!/bin/bash
# tensor-core attention backward: gradient parity tests, then the training-step timing with both attention-backward versions
o=gpurun_out; mkdir -p $o
timeout 900 python -m pytest tests/test_gpu_train.py tests/test_gpu_parity.py -m gpu -x -q -k "train or schedules or bptt or decoder_b" > $o/bwd_tc_tests.log 2>&1; echo "pytest rc=$?" >> $o/bwd_tc_tests.log
tail -15 $o/bwd_tc_tests.log
timeout 600 python bench.py --train-only --steps 5 --warmup 3 > $o/bwd_tc_train.json 2> $o/bwd_tc_train.err; tail -2 $o/bwd_tc_train.err
T2B200_BWD_ATT=cuda timeout 600 python bench.py --train-only --steps 5 --warmup 3 > $o/bwd_cuda_train.json 2>> $o/bwd_tc_train.err
python - <<'PY'
import json
for f in ("gpurun_out/bwd_tc_train.json","gpurun_out/bwd_cuda_train.json"):
    try:
        d=json.load(open(f)); t=d.get("train",d)
        print(f, t["ms_per_step"], t["phase_ms"], t["roofline"]["kernel_ms"]); print(t["decoder_bwd_phase_us_per_frame"])
    except Exception as e: print(f, "ERR", e)
PY
