#!/bin/bash
# the allocation-free entry point alone in a fresh process, then the full GPU suite (which ends with it), then the evidence set
o=gpurun_out; tag=${1:-r02_v3}
mkdir -p $o
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rs > $o/${tag}_no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_no_alloc_alone.log
tail -4 $o/${tag}_no_alloc_alone.log
bash tools/r02_call2.sh $tag
