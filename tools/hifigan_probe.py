"""HiFi-GAN generator forward alone (profiling aid): python tools/hifigan_probe.py [B] [T]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import hifigan  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
T = int(sys.argv[2]) if len(sys.argv) > 2 else 800
print(json.dumps(hifigan.bench(torch.device("cuda:0"), B, T, steps=2)))
