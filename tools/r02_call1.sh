#!/bin/bash
# round 2, GPU call 1: full GPU test suite, both bench arms, ncu metric pass over the tensor-core kernels off the recurrence
o=gpurun_out; tag=${1:-r02_v1}
mkdir -p $o
timeout 1200 python -m pytest tests -m gpu -q -rs > $o/${tag}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_gpu_tests.log
tail -5 $o/${tag}_gpu_tests.log
timeout 900 python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>> $o/${tag}_bench.err
head -c 600 $o/${tag}_bench.json; echo; tail -3 $o/${tag}_bench.err
M=gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_uniform.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__throughput.avg.pct_of_peak_sustained_elapsed,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active
python tools/kernel_probe.py > $o/${tag}_kernel_probe.log 2>&1 && \
ncu --metrics $M --clock-control none -k regex:'conv_tc|gemm_nt|bilstm' --csv --log-file $o/${tag}_tc_kernels.csv python tools/kernel_probe.py > $o/${tag}_ncu_probe.log 2>&1
tail -2 $o/${tag}_kernel_probe.log; wc -l $o/${tag}_tc_kernels.csv
