# scratch script of the gpurun experiments (profiles/r01_v12_experiments/README.md lists what each run compared)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
python tools/train_probe.py 4 2>&1 | tail -3 | cut -c1-330
python tools/phase_probe.py 2>&1 | tail -1
