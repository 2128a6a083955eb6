timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
python tools/phase_probe.py 2>&1 | tail -1
