"""Generates tests/golden/collate_golden.pt from the REFERENCE's own `_collate` (datasets/tts_dataloader.py:8-35) and the gate
construction of datasets/tts_dataset.py:211-214.  Run in the build container (needs /root/reference):
    python tests/golden/make_collate_golden.py
"""
import os
import sys

import torch

sys.path.insert(0, "/root/reference")
from datasets.tts_dataloader import _collate  # noqa: E402  (imports torch only)


def items(seed=11, B=5, n_mels=80):
    g = torch.Generator().manual_seed(seed)
    out = []
    for b in range(B):
        T = int(torch.randint(3, 23, (1,), generator=g))
        L = int(torch.randint(2, 17, (1,), generator=g))
        mel = torch.randn(T, n_mels, generator=g)
        gate = torch.ones(len(mel), 1)   # datasets/tts_dataset.py:211-214
        gate[-1] = 0.0
        data = {"chars_idx": torch.randint(1, 40, (L,), generator=g, dtype=torch.int64), "mel_spectrogram": mel, "gate": gate,
                "features": torch.rand(5, generator=g).unsqueeze(0)}
        lens = {"chars_idx_len": torch.tensor([L], dtype=torch.int64), "mel_spectrogram_len": torch.IntTensor([T]),
                "gate_len": torch.IntTensor([T])}
        out.append((data, lens, {"text": f"utt{b}"}))
    return out


if __name__ == "__main__":
    it = items()
    d, n, e = _collate(it)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "collate_golden.pt")
    torch.save({"items": it, "data": d, "len": n, "extra": e}, path)
    print(path, {k: tuple(v.shape) for k, v in d.items()}, {k: v.tolist() for k, v in n.items()})
