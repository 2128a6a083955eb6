#!/bin/bash
# Round evidence on one B200: bench (both arms), ncu launch list of the same command, ncu --set full of the decoder kernel.
# usage: tools/evidence.sh <tag>     (writes gpurun_out/<tag>_*)
tag=${1:-r02_vX}
o=gpurun_out
python bench.py > $o/${tag}_bench.json 2> $o/${tag}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > $o/${tag}_bench_reference_arm.json 2>> $o/${tag}_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/${tag}_launches.csv python bench.py --steps 1 --warmup 1 > $o/${tag}_ncu_bench.log 2>&1
PROBE_T=200 ncu --set full --clock-control none --import-source on -k regex:decoder_ar_kernel -s 3 -c 1 -f -o $o/${tag}_decoder_fwd python tools/phase_probe.py > $o/${tag}_ncu_full.log 2>&1
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page raw --csv > $o/${tag}_decoder_fwd_full_raw.csv 2>/dev/null
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page details > $o/${tag}_decoder_fwd_details.txt 2>/dev/null
ncu -i $o/${tag}_decoder_fwd.ncu-rep --page source --csv --print-source cuda,sass > $o/${tag}_decoder_fwd_source.csv 2>/dev/null
python tools/ncu_lines.py $o/${tag}_decoder_fwd_source.csv 40 > $o/${tag}_decoder_fwd_top_stalls.txt 2>&1
python tools/ncu_summary.py $o/${tag}_launches.csv > $o/${tag}_launch_summary.txt 2>&1
rm -f $o/${tag}_decoder_fwd_source.csv
head -c 400 $o/${tag}_bench.json; echo; tail -3 $o/${tag}_bench.err; head -5 $o/${tag}_launch_summary.txt
