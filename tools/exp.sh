python tools/phase_probe.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_train.py tests/test_tts_model.py -m gpu -x -q 2>&1 | tail -3
T2B200_NO_PROJ_PART=1 python tools/phase_probe.py 2>&1 | tail -1
python tools/phase_probe.py 2>&1 | tail -1
