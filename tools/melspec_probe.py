"""Log-mel extraction probe: C3-sized batch (64 x 800 frames), kernel time by CUDA events."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import synth  # noqa: E402
from tacotron2_b200.melspec import TacotronMelSpectrogram  # noqa: E402

dev = torch.device("cuda:0")
B, Tm = 64, 800
S = (Tm - 1) * 256
wav = torch.from_numpy(np.stack([synth.synth_wav(S, seed=i % 4) for i in range(4)])).repeat(B // 4, 1).to(dev)
lens = torch.full((B,), S, dtype=torch.int64, device=dev)
tr = TacotronMelSpectrogram().to(dev)
for _ in range(3):
    out, n = tr.batch(wav, lens)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    out, n = tr.batch(wav, lens)
e1.record()
e1.synchronize()
print("melspec ms %.4f  frames %d" % (e0.elapsed_time(e1) / 10, int(n.sum())), flush=True)
