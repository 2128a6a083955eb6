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


def test_trim_oracle_properties_cpu():
    """The CPU restatement of librosa.effects.trim (parity unpinned vs librosa, absent here): silence around a burst is removed on
    512-sample frame boundaries, an all-quiet signal trims to nothing, a loud signal is kept whole."""
    import numpy as np
    from oracle import trim_oracle as T
    y = np.zeros(22050, np.float32)
    y[5000:9000] = 0.5 * np.sin(np.arange(4000) * 0.1).astype(np.float32)
    out, (s, e) = T.trim(y, top_db=40.0, frame_length=2048, hop_length=512)
    assert s % 512 == 0 and s <= 5000 < s + 2048 and 9000 <= e <= 9000 + 2048 and out.shape[0] == e - s
    assert T.trim(np.zeros(4000, np.float32))[1] == (0, 4000)          # every frame is at the amin floor = the max: nothing is below -top_db
    assert T.trim(np.ones(3000, np.float32) * 0.3)[1] == (0, 3000)


@pytest.mark.gpu
def test_gpu_trim_matches_oracle():
    import numpy as np
    from oracle import trim_oracle as T
    from tacotron2_b200 import synth
    from tacotron2_b200.collate import trim_bounds
    dev = torch.device("cuda:0")
    wavs = []
    for i, (n, lead, tail) in enumerate(((22050, 3000, 4000), (9000, 0, 2500), (5120, 700, 0), (2048, 100, 100), (40000, 12345, 777))):
        w = synth.synth_wav(n, seed=i)
        w = np.concatenate([np.zeros(lead, np.float32) + 1e-6 * np.sin(np.arange(lead)), w, np.zeros(tail, np.float32)]).astype(np.float32)
        wavs.append(w)
    L = max(len(w) for w in wavs)
    pad = torch.zeros(len(wavs), L)
    for i, w in enumerate(wavs):
        pad[i, : len(w)] = torch.from_numpy(w)
    lens = torch.tensor([len(w) for w in wavs])
    for top_db, fl in ((60.0, 2048), (30.0, 1024), (80.0, 512)):
        s, e = trim_bounds(pad.to(dev), lens, top_db, fl)
        want = [T.trim(w, top_db=top_db, frame_length=fl)[1] for w in wavs]
        assert list(zip(s.tolist(), e.tolist())) == want, (top_db, fl, s.tolist(), e.tolist(), want)


@pytest.mark.gpu
def test_assemble_items_with_trim_and_mel_cache(tmp_path):
    """Trim + mel on the device + the reference's .pt cache format: the first pass computes and writes the cache, the second pass reads
    it (and a file written the way the reference writes it is read back the same way)."""
    import numpy as np
    from oracle import melspec_oracle as MO
    from oracle import trim_oracle as T
    from tacotron2_b200 import synth
    from tacotron2_b200.collate import MelCache, assemble_items
    dev = torch.device("cuda:0")
    names = ["spk/a.wav", "spk/b.wav", "c.wav"]
    wav = {n: torch.from_numpy(np.concatenate([np.zeros(1500, np.float32), synth.synth_wav(6000 + 1700 * i, seed=i)])) for i, n in enumerate(names)}
    chars = [torch.arange(1, 5 + i) for i in range(3)]
    cache = MelCache(str(tmp_path / "mel"))
    torch.save(torch.full((7, 80), -3.0), cache.path("c.wav"))     # an item the reference cached earlier
    d1, l1 = assemble_items(names, lambda f: wav[f], chars, dev, cache, trim=True, trim_top_db=40.0, trim_frame_length=1024, silence=256)
    assert cache.has("spk/a.wav") and cache.path("spk/a.wav").endswith("spk_a.wav.pt")
    assert l1["mel_spectrogram_len"].tolist()[2] == 7 and float(d1["mel_spectrogram"][2, :7].mean()) == -3.0
    for i, n in enumerate(names[:2]):
        y, _ = T.trim(wav[n].numpy(), top_db=40.0, frame_length=1024)
        want = MO.log_mel(np.concatenate([y, np.zeros(256, np.float32)]))
        got = d1["mel_spectrogram"][i, : l1["mel_spectrogram_len"][i]].cpu().numpy()
        assert got.shape == want.shape and float(np.abs(got - want).max()) <= 1e-3
        assert torch.equal(torch.load(cache.path(n)), d1["mel_spectrogram"][i, : l1["mel_spectrogram_len"][i]].cpu())
        assert float(d1["gate"][i, l1["gate_len"][i] - 1]) == 0.0 and float(d1["gate"][i, 0]) == 1.0
    d2, l2 = assemble_items(names, lambda f: (_ for _ in ()).throw(AssertionError("cache miss")), chars, dev, cache)
    assert torch.equal(d1["mel_spectrogram"], d2["mel_spectrogram"]) and torch.equal(l1["mel_spectrogram_len"], l2["mel_spectrogram_len"])
