#!/bin/bash
# full GPU suite with the pipelined AR kernel as default + evidence (bench both arms, launch list, ncu --set full of decoder_ar_kernel)
o=gpurun_out; tag=${1:-r02_v2}
timeout 900 python -m pytest tests -m gpu -q -rs > $o/${tag}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_gpu_tests.log
tail -4 $o/${tag}_gpu_tests.log
bash tools/evidence.sh $tag
