"""Log-mel extraction: oracle vs torchaudio golden (CPU), CUDA kernel vs both (GPU).

Tolerance: max-abs log-mel error <= 1e-3 (0.1 % of the mel magnitude; fp32 FFT round-off is ~1e-5), frame counts and
zero padding exact."""
import os

import numpy as np
import pytest
import torch

from oracle import melspec_oracle as MO
from tacotron2_b200 import synth

G = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "melspec.npz")))
TOL = 1e-3


def test_oracle_filterbank_matches_torchaudio():
    assert np.abs(MO.filterbank() - G["fb"]).max() <= 1e-6


def test_host_filterbank_matches_torchaudio():
    from tacotron2_b200.melspec import slaney_filterbank
    assert np.abs(slaney_filterbank(22050, 80).T - G["fb"]).max() <= 1e-6


@pytest.mark.parametrize("i", range(len(synth.MEL_GOLDEN_LENGTHS)))
def test_oracle_matches_torchaudio(i):
    S = synth.MEL_GOLDEN_LENGTHS[i]
    m = MO.log_mel(synth.synth_wav(S, seed=i))
    assert m.shape == G[f"mel{i}"].shape == (1 + S // 256, 80)
    assert np.abs(m - G[f"mel{i}"]).max() <= 5e-5


@pytest.mark.gpu
def test_gpu_single_utterance_golden():
    from tacotron2_b200.melspec import TacotronMelSpectrogram
    tr = TacotronMelSpectrogram(n_mels=80, sample_rate=22050).cuda()
    for i, S in enumerate(synth.MEL_GOLDEN_LENGTHS):
        wav = torch.from_numpy(synth.synth_wav(S, seed=i)).cuda()
        got = tr(wav, id=str(i)).cpu().numpy()  # call form of datasets/tts_dataset.py:204
        assert got.shape == G[f"mel{i}"].shape
        assert np.abs(got - G[f"mel{i}"]).max() <= TOL, (i, np.abs(got - G[f"mel{i}"]).max())


@pytest.mark.gpu
def test_gpu_ragged_batch_vs_oracle():
    from tacotron2_b200.melspec import TacotronMelSpectrogram
    tr = TacotronMelSpectrogram().cuda()
    lens = [40000, 22050, 12345, 600, 256 * 64, 256 * 64 - 1, 33333]
    S = max(lens)
    wavs = np.zeros((len(lens), S), np.float32)
    for b, n in enumerate(lens):
        wavs[b, :n] = synth.synth_wav(n, seed=100 + b)
    out, nf = tr.batch(torch.from_numpy(wavs).cuda(), torch.tensor(lens))
    out, nf = out.cpu().numpy(), nf.cpu().numpy()
    assert out.shape == (len(lens), 1 + S // 256, 80)
    for b, n in enumerate(lens):
        want = MO.log_mel(wavs[b, :n])
        assert nf[b] == want.shape[0] == 1 + n // 256
        assert np.abs(out[b, :nf[b]] - want).max() <= TOL, (b, np.abs(out[b, :nf[b]] - want).max())
        assert np.all(out[b, nf[b]:] == 0.0)  # pad_sequence zero padding (datasets/tts_dataloader.py:28)


@pytest.mark.gpu
def test_gpu_properties_full_size():
    """Size-independent properties at bench size (64 x 10 s): bit-determinism, batch independence, a pure tone lands in
    the right filter, silence hits the log floor exactly."""
    from tacotron2_b200.melspec import TacotronMelSpectrogram
    tr = TacotronMelSpectrogram().cuda()
    B, S = 64, 220500
    g = torch.Generator().manual_seed(5)
    wavs = (torch.randn(B, S, generator=g) * 0.1).cuda()
    t = torch.arange(S, device="cuda") / 22050.0
    wavs[0] = 0.5 * torch.sin(2 * np.pi * 1000.0 * t)
    wavs[1] = 0.0
    lens = torch.full((B,), S)
    a, n1 = tr.batch(wavs, lens)
    b, _ = tr.batch(wavs, lens)
    assert torch.equal(a, b)
    assert int(n1[0]) == 1 + S // 256
    one, _ = tr.batch(wavs[5:6], lens[5:6])
    assert torch.equal(one[0], a[5])
    fb = torch.from_numpy(G["fb"])
    k1000 = int(round(1000.0 / (22050 / 1024)))
    assert int(a[0, 100].argmax()) == int(fb[k1000].argmax())
    assert float((a[1] - float(np.log(1e-5))).abs().max()) <= 1e-5
