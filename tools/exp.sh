for v in nb2 nb3 nb6; do cp tools/ab/$v.so tacotron2_b200/libt2b200.so; echo $v; python tools/melspec_probe.py 2>&1 | tail -1; done
