#!/bin/bash
# Round evidence on one B200 (second form): the allocation-free entry point alone, the full GPU suite, bench (both arms, no profiler),
# the ncu launch list of the same command, ncu --set full of decoder_ar_kernel (200-frame launch of the C2 batch) and of
# decoder_bwd_kernel (one C3 training step), DRAM bytes per frame of both -> ncu_traffic.json.   usage: tools/evidence2.sh <tag>
tag=${1:-r02_vX}
o=gpurun_out
mkdir -p $o
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -rs > $o/${tag}_no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_no_alloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -rs > $o/${tag}_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $o/${tag}_gpu_tests.log
tail -3 $o/${tag}_gpu_tests.log
python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>> $o/${tag}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/${tag}_launches.csv python bench.py --steps 1 --warmup 1 > $o/${tag}_ncu_bench.log 2>&1
python tools/ncu_summary.py $o/${tag}_launches.csv > $o/${tag}_launch_summary.txt 2>&1
PROBE_T=200 ncu --set full --clock-control none --import-source on -k regex:decoder_ar_kernel -s 3 -c 1 -f -o $o/${tag}_decoder_fwd python tools/phase_probe.py > $o/${tag}_ncu_full.log 2>&1
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page raw --csv > $o/${tag}_decoder_fwd_full_raw.csv 2>/dev/null
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page details > $o/${tag}_decoder_fwd_details.txt 2>/dev/null
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page source --csv --print-source cuda,sass > $o/${tag}_decoder_fwd_source.csv 2>/dev/null
python tools/ncu_lines.py $o/${tag}_decoder_fwd_source.csv 40 > $o/${tag}_decoder_fwd_top_stalls.txt 2>&1
rm -f $o/${tag}_decoder_fwd_source.csv $o/${tag}_decoder_fwd.ncu-rep
ncu --set full --clock-control none --import-source on -k regex:decoder_bwd_kernel -s 1 -c 1 -f -o $o/${tag}_decoder_bwd python tools/train_probe.py 3 > $o/${tag}_ncu_bwd.log 2>&1
ncu -i $o/${tag}_decoder_bwd.ncu-rep --page raw --csv > $o/${tag}_decoder_bwd_full_raw.csv 2>/dev/null
ncu -i $o/${tag}_decoder_bwd.ncu-rep --page source --csv --print-source cuda,sass > $o/${tag}_decoder_bwd_source.csv 2>/dev/null
python tools/ncu_lines.py $o/${tag}_decoder_bwd_source.csv 40 > $o/${tag}_decoder_bwd_top_stalls.txt 2>&1
rm -f $o/${tag}_decoder_bwd_source.csv $o/${tag}_decoder_bwd.ncu-rep
python tools/ncu_traffic.py $o/ncu_traffic.json \
  decoder_ar_kernel $o/${tag}_decoder_fwd_full_raw.csv 200 "profiles/${tag}_decoder_fwd_full_raw.csv (ncu --set full of decoder_ar_kernel, 200-frame launch of the C2 batch; per frame x frames)" \
  decoder_bwd_kernel $o/${tag}_decoder_bwd_full_raw.csv 800 "profiles/${tag}_decoder_bwd_full_raw.csv (ncu --set full of decoder_bwd_kernel, one C3 training step of 800 frames; per frame x frames)" > $o/${tag}_ncu_traffic.log 2>&1
head -c 600 $o/${tag}_bench.json; echo; tail -3 $o/${tag}_bench.err; head -5 $o/${tag}_launch_summary.txt; cat $o/${tag}_ncu_traffic.log | head -20
