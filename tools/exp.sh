# scratch script of the gpurun experiments (profiles/r01_v12_experiments/README.md lists what each run compared)
python tools/phase_probe.py 2>&1 | tail -1
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
