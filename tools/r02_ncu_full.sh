#!/bin/bash
# ncu --set full of decoder_ar_kernel (200-frame launch of the C2 batch) for library builds LIBS="name ..." ("tree" = working tree)
o=gpurun_out
for n in $LIBS; do
  lib=build/libt2b200_$n.so; [ $n = tree ] && lib=tacotron2_b200/libt2b200.so
  T2B200_LIB=$lib PROBE_T=200 ncu --set full --clock-control none --import-source on -k regex:decoder_ar_kernel -s 3 -c 1 -f -o $o/full_$n python tools/phase_probe.py > $o/full_$n.log 2>&1
  ncu -i $o/full_$n.ncu-rep --page raw --csv > $o/${TAG:-r02_v5}_full_raw_$n.csv 2>/dev/null
  ncu -i $o/full_$n.ncu-rep --page source --csv --print-source cuda,sass > $o/${TAG:-r02_v5}_source_$n.csv 2>/dev/null
  python tools/ncu_lines.py $o/${TAG:-r02_v5}_source_$n.csv 70 > $o/${TAG:-r02_v5}_stalls_$n.txt 2>&1
  gzip -f $o/${TAG:-r02_v5}_source_$n.csv
  rm -f $o/full_$n.ncu-rep
  tail -1 $o/full_$n.log
done
