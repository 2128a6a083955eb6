timeout 900 python -m pytest tests/test_gpu_train.py tests/test_tts_model.py -m gpu -x -q 2>&1 | tail -3
python tools/train_probe.py 4 2>&1 | tail -12
