#!/bin/bash
# AR kernel A/B on one box: build/libt2b200_base.so (previous revision) against the working tree; experiment bits of the new build
o=gpurun_out
( echo "== base"; T2B200_LIB=build/libt2b200_base.so python tools/ar_probe.py 64 200 400
  echo "== new"; python tools/ar_probe.py 64 200 400
  for b in $BITS; do echo "== new bits $b"; T2B200_DEBUG_BARRIER=$b python tools/ar_probe.py 64 200 400; done
) > $o/${TAG:-r02_v5}_ar_ab.log 2>&1
grep -v "^$" $o/${TAG:-r02_v5}_ar_ab.log | cut -c1-700
