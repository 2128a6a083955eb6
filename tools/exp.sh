run() { cp tools/ab/$1.so tacotron2_b200/libt2b200.so; shift; echo "== $*"; env "$@" python tools/phase_probe.py 2>&1 | tail -1; }
run eu4 X=eu4
run eu2 X=eu2
run eu8 X=eu8
run eu4 X=eu4
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
