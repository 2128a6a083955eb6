#!/bin/bash
# same-box A/B over library builds: LIBS="name[:bits] ..." (build/libt2b200_<name>.so; "tree" = the working tree's library; bits = T2B200_DEBUG_BARRIER)
o=gpurun_out
( for e in $LIBS; do
    n=${e%%:*}; bits=0; [ "$e" != "$n" ] && bits=${e##*:}
    echo "== $n bits $bits"
    lib=build/libt2b200_$n.so; [ $n = tree ] && lib=tacotron2_b200/libt2b200.so
    T2B200_DEBUG_BARRIER=$bits T2B200_LIB=$lib python tools/ar_probe.py 64 200 400
  done ) > $o/${TAG:-r02_v5}_ar_ab.log 2>&1
grep -v "^$" $o/${TAG:-r02_v5}_ar_ab.log | cut -c1-700
