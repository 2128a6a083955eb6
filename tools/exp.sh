python tools/phase_probe.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_train.py tests/test_collate.py -m gpu -x -q 2>&1 | tail -3
python tools/phase_probe.py 2>&1 | tail -1
