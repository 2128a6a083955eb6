#!/bin/bash
# warp-stall samples per source line of decoder_ar_kernel for several library builds (LIBS="name ...", "tree" = working tree)
o=gpurun_out
for n in $LIBS; do
  lib=build/libt2b200_$n.so; [ $n = tree ] && lib=tacotron2_b200/libt2b200.so
  T2B200_LIB=$lib PROBE_T=200 ncu --section SourceCounters --section WarpStateStats --section SchedulerStats --clock-control none --import-source on -k regex:decoder_ar_kernel -s 3 -c 1 -f -o $o/src_$n python tools/phase_probe.py > $o/src_$n.log 2>&1
  ncu -i $o/src_$n.ncu-rep --page source --csv --print-source cuda,sass > $o/src_$n.csv 2>/dev/null
  python tools/ncu_lines.py $o/src_$n.csv 70 > $o/${TAG:-r02_v5}_stalls_$n.txt 2>&1
  ncu -i $o/src_$n.ncu-rep --page details > $o/${TAG:-r02_v5}_details_$n.txt 2>/dev/null
  rm -f $o/src_$n.csv $o/src_$n.ncu-rep
  head -3 $o/${TAG:-r02_v5}_stalls_$n.txt; tail -2 $o/src_$n.log
done
