"""Log-mel feature extraction on the GPU.

Mirrors the object the reference builds at ``datasets/tts_dataset.py:166-168`` and calls at
``:204`` -- ``TacotronMelSpectrogram(n_mels=80, sample_rate=22050)(wav, id=str(i)) -> (T, n_mels)``.
The reference takes that class from ``speech_utils`` (git dependency, un-vendored, unpinned:
``requirements.txt:12``); its parameters are stated in-repo at
``datasets/prosody_dataset.py:39-50,66-67`` (torchaudio ``MelSpectrogram(n_fft 1024, win 1024,
hop 256, f_min 0, f_max 8000, power 1, mel_scale/norm "slaney")`` then ``log(clamp(min=1e-5)).T``)
and confirmed by the inverse at ``run/say.py:161-171``.  Parity is therefore "unpinned" with
respect to speech_utils itself and pinned against torchaudio (tests/golden/melspec.npz).

The filterbank / window tables are parameter preparation (host, float64 numpy, once); the
per-sample arithmetic -- reflect padding, windowing, FFT, magnitude, filterbank, log -- runs in
``t2_melspec`` (csrc/melspec.cu).  CUDA tensors only; no fallback.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib

N_FFT, WIN, HOP, F_MIN, F_MAX, LOG_FLOOR = 1024, 1024, 256, 0.0, 8000.0, 1e-5


def _hz_to_mel_slaney(f):
    f = np.asarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    with np.errstate(divide="ignore", invalid="ignore"):
        log_t = min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep
    return np.where(f >= min_log_hz, log_t, mels)


def _mel_to_hz_slaney(m):
    m = np.asarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    return np.where(m >= min_log_mel, min_log_hz * np.exp(logstep * (m - min_log_mel)), f_sp * m)


def slaney_filterbank(sample_rate: int, n_mels: int, n_fft: int = N_FFT, f_min: float = F_MIN, f_max: float = F_MAX):
    """(n_mels, n_fft//2+1) float32 triangular filters, slaney scale and area normalisation
    (what ``torchaudio.functional.melscale_fbanks(..., norm="slaney", mel_scale="slaney")`` builds, transposed)."""
    n_freqs = n_fft // 2 + 1
    all_freqs = np.linspace(0.0, sample_rate // 2, n_freqs)
    m_pts = np.linspace(_hz_to_mel_slaney(f_min), _hz_to_mel_slaney(f_max), n_mels + 2)
    f_pts = _mel_to_hz_slaney(m_pts)
    f_diff = f_pts[1:] - f_pts[:-1]
    slopes = f_pts[None, :] - all_freqs[:, None]
    down = -slopes[:, :-2] / f_diff[:-1]
    up = slopes[:, 2:] / f_diff[1:]
    fb = np.maximum(0.0, np.minimum(down, up))
    fb = fb * (2.0 / (f_pts[2:n_mels + 2] - f_pts[:n_mels]))[None, :]
    return np.ascontiguousarray(fb.T.astype(np.float32))


class TacotronMelSpectrogram(torch.nn.Module):
    """Call-compatible with the object at datasets/tts_dataset.py:166-168: ``__call__(wav, id=None) -> (T, n_mels)``.

    ``batch(wavs, lens)`` is the batched form a GPU-side collate uses: (B,S) zero-padded waveforms ->
    (B, T_max, n_mels) zero-padded log-mels + (B,) frame counts, i.e. ``mel_spectrogram`` /
    ``mel_spectrogram_len`` of datasets/tts_dataloader.py:26-33.
    """

    def __init__(self, n_mels: int = 80, sample_rate: int = 22050):
        super().__init__()
        self.n_mels, self.sample_rate = n_mels, sample_rate
        fb = slaney_filterbank(sample_rate, n_mels)
        rng = np.zeros((n_mels, 2), np.int32)
        for m in range(n_mels):
            nz = np.nonzero(fb[m])[0]
            if nz.size:
                rng[m] = (nz[0], nz[-1] + 1)
        n = np.arange(WIN, dtype=np.float64)
        self.register_buffer("fb", torch.from_numpy(fb), persistent=False)
        self.register_buffer("fb_range", torch.from_numpy(rng), persistent=False)
        self.register_buffer("window", torch.from_numpy((0.5 - 0.5 * np.cos(2 * np.pi * n / WIN)).astype(np.float32)), persistent=False)

    def batch(self, wavs: torch.Tensor, lens: torch.Tensor):
        if wavs.device.type != "cuda":
            raise RuntimeError("tacotron2_b200 runs on CUDA (sm_100a) tensors only; there is no CPU fallback")
        lib = _lib.load()
        dev = wavs.device
        if self.fb.device != dev:
            self.to(dev)
        wavs = wavs.to(torch.float32).contiguous()
        lens = lens.to(device=dev, dtype=torch.int64).contiguous()
        B, S = wavs.shape
        t_max = 1 + S // HOP
        out = torch.empty(B, t_max, self.n_mels, device=dev)
        n_frames = torch.empty(B, dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.t2_melspec(_lib.ptr(wavs), _lib.ptr(lens), B, S, _lib.ptr(self.window), _lib.ptr(self.fb),
                                      _lib.ptr(self.fb_range), N_FFT, HOP, self.n_mels, _lib.ptr(out), t_max, _lib.ptr(n_frames),
                                      st), "t2_melspec")
        return out, n_frames

    def forward(self, wav: torch.Tensor, id=None):  # noqa: A002  (the reference passes id=str(i))
        out, _ = self.batch(wav.reshape(1, -1), torch.tensor([wav.numel()]))
        return out[0]
