"""Batch assembly: the oracle against the reference's `_collate` (golden fixture generated from the
reference, plus the reference itself when /root/reference is present), and -- on the GPU -- `t2_collate` through the host mirror
against the oracle, bit for bit."""
import os
import sys

import pytest
import torch

from oracle import collate_oracle as O
from tacotron2_b200 import collate as Cc

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden", "collate_golden.pt")


def _same(a, b):
    assert set(a) == set(b), set(a) ^ set(b)
    for k in a:
        x, y = a[k].cpu(), b[k].cpu()
        assert x.dtype == y.dtype and x.shape == y.shape, (k, x.dtype, y.dtype, x.shape, y.shape)
        assert torch.equal(x, y), k


def test_oracle_matches_golden():
    g = torch.load(GOLD)
    d, n, e = O.collate(g["items"])
    _same(d, g["data"])
    _same(n, g["len"])
    assert e == g["extra"]
    for b, (item, _, _) in enumerate(g["items"]):
        T = item["mel_spectrogram"].shape[0]
        assert torch.equal(O.gate_target(T), item["gate"])


@pytest.mark.skipif(not os.path.isdir("/root/reference/datasets"), reason="reference tree not present")
def test_oracle_matches_reference_collate():
    sys.path.insert(0, "/root/reference")
    try:
        from datasets.tts_dataloader import _collate
    finally:
        sys.path.pop(0)
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import make_collate_golden as M
    for seed, B in ((1, 2), (2, 7), (3, 1)):
        items = M.items(seed=seed, B=B)
        d0, n0, e0 = _collate(items)
        d1, n1, e1 = O.collate(items)
        _same(d1, d0)
        if B > 1:   # (torch.stack(...).squeeze(1) of a single (1,) tensor keeps shape (1,) either way)
            _same(n1, n0)
        assert e1 == e0


def test_no_cpu_fallback():
    g = torch.load(GOLD)
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        Cc.collate(g["items"], device="cpu")


def _ragged(seed, B, tmax, lmax, n_mels=80):
    g = torch.Generator().manual_seed(seed)
    items = []
    for b in range(B):
        T = int(torch.randint(1, tmax + 1, (1,), generator=g))
        L = int(torch.randint(1, lmax + 1, (1,), generator=g))
        d = {"chars_idx": torch.randint(1, 40, (L,), generator=g, dtype=torch.int64), "mel_spectrogram": torch.randn(T, n_mels, generator=g),
             "gate": O.gate_target(T)}
        items.append((d, {"chars_idx_len": torch.tensor([L]), "mel_spectrogram_len": torch.IntTensor([T]), "gate_len": torch.IntTensor([T])},
                      {"text": str(b)}))
    return items


@pytest.mark.gpu
def test_gpu_collate_matches_golden():
    g = torch.load(GOLD)
    d, n, e = Cc.collate(g["items"], "cuda:0")
    torch.cuda.synchronize()
    assert all(v.is_cuda for v in d.values())
    _same(d, g["data"])
    _same(n, g["len"])
    assert e == g["extra"]


@pytest.mark.gpu
@pytest.mark.parametrize("B,tmax,lmax", [(1, 5, 3), (3, 1, 1), (8, 600, 150), (64, 800, 200)])
def test_gpu_collate_matches_oracle(B, tmax, lmax):
    items = _ragged(100 + B, B, tmax, lmax)
    want, wn, _ = O.collate(items)
    got, gn, _ = Cc.collate(items, "cuda:0")
    torch.cuda.synchronize()
    _same(got, want)
    if B > 1:
        _same(gn, wn)
    # gate generated on the device when the items do not carry it (datasets/tts_dataset.py:211-214)
    bare = [({k: v for k, v in d.items() if k != "gate"}, n, e) for d, n, e in items]
    got2, _, _ = Cc.collate(bare, "cuda:0")
    torch.cuda.synchronize()
    _same(got2, want)


@pytest.mark.gpu
def test_gpu_collate_feeds_the_model():
    """The assembled batch has the layout Tacotron2.forward expects: one teacher-forced forward runs on it."""
    from tacotron2_b200 import synth
    from tacotron2_b200.model import Tacotron2
    items = _ragged(5, 4, 12, 9)
    d, n, _ = Cc.collate(items, "cuda:0")
    m = Tacotron2(**synth.VANILLA.kwargs()).to("cuda:0").eval()
    m.load_state_dict(synth.synth_state_dict(synth.VANILLA, seed=0))
    with torch.no_grad():
        mels, post, gates, al = m(d["chars_idx"], n["chars_idx_len"], True, d["mel_spectrogram"], n["mel_spectrogram_len"])
    assert mels.shape == d["mel_spectrogram"].shape and gates.shape == d["gate"].shape and torch.isfinite(post).all()


@pytest.mark.gpu
def test_gpu_assemble_from_wavs():
    """wav -> (silence pad) -> log-mel -> padded batch + gate targets, all on the GPU (datasets/tts_dataset.py:201-214 +
    datasets/tts_dataloader.py:8-35), against the per-item CPU oracles."""
    import numpy as np
    from oracle import melspec_oracle as MO
    from tacotron2_b200 import synth
    lens, silence = [9000, 22050, 4100, 256 * 20, 777], 300
    wavs = [torch.from_numpy(synth.synth_wav(n, seed=50 + i)) for i, n in enumerate(lens)]
    g = torch.Generator().manual_seed(5)
    chars = [torch.randint(1, 40, (int(torch.randint(2, 30, (1,), generator=g)),), generator=g, dtype=torch.int64) for _ in lens]
    data, ln = Cc.assemble_from_wavs(wavs, chars, "cuda:0", silence=silence)
    torch.cuda.synchronize()
    items = []
    for w, c in zip(wavs, chars):
        mel = torch.from_numpy(MO.log_mel(np.concatenate([w.numpy(), np.zeros(silence, np.float32)])))
        items.append(({"chars_idx": c, "mel_spectrogram": mel, "gate": O.gate_target(mel.shape[0])},
                      {"chars_idx_len": torch.tensor([c.numel()]), "mel_spectrogram_len": torch.IntTensor([mel.shape[0]]),
                       "gate_len": torch.IntTensor([mel.shape[0]])}, {}))
    want, wl, _ = O.collate(items)
    assert torch.equal(data["chars_idx"].cpu(), want["chars_idx"]) and torch.equal(data["gate"].cpu(), want["gate"])
    assert data["mel_spectrogram"].shape == want["mel_spectrogram"].shape
    assert float((data["mel_spectrogram"].cpu() - want["mel_spectrogram"]).abs().max()) <= 1e-3   # the melspec parity bar
    for k in wl:
        assert ln[k].dtype == wl[k].dtype and torch.equal(ln[k].cpu(), wl[k]), k
