"""CPU oracle for log-mel extraction.  TEST INFRASTRUCTURE ONLY (tests/, smoke(), bench.py cpu legs).

Restates, in float64/float32 numpy, what the reference computes at datasets/tts_dataset.py:204 through
``TacotronMelSpectrogram`` -- an un-vendored, unpinned dependency (speech_utils, requirements.txt:12) whose
parameters the reference states at datasets/prosody_dataset.py:39-50,66-67 and run/say.py:161-171:
torchaudio ``MelSpectrogram(sr 22050, n_fft 1024, win 1024, hop 256, f_min 0, f_max 8000, n_mels 80, power 1,
mel_scale="slaney", norm="slaney")`` (torch.stft center=True, reflect pad, periodic Hann), then
``log(clamp(min=1e-5))`` transposed to (T, n_mels).

Parity status: UNPINNED against speech_utils (absent, no reference test or golden vector covers it);
pinned against torchaudio 2.11 -- the class the in-repo statement instantiates -- through
tests/golden/melspec.npz (tests/golden/make_golden_mel.py).
"""
import numpy as np

N_FFT, HOP = 1024, 256


def hz_to_mel(f):
    f = np.asarray(f, np.float64)
    lin = f / (200.0 / 3)
    log = 15.0 + np.log(np.maximum(f, 1e-300) / 1000.0) / (np.log(6.4) / 27.0)
    return np.where(f >= 1000.0, log, lin)


def mel_to_hz(m):
    m = np.asarray(m, np.float64)
    return np.where(m >= 15.0, 1000.0 * np.exp((np.log(6.4) / 27.0) * (m - 15.0)), (200.0 / 3) * m)


def filterbank(sample_rate=22050, n_mels=80, f_min=0.0, f_max=8000.0):
    """torchaudio.functional.melscale_fbanks(513, f_min, f_max, n_mels, sr, norm='slaney', mel_scale='slaney') -> (513, n_mels)."""
    freqs = np.linspace(0, sample_rate // 2, N_FFT // 2 + 1)
    f_pts = mel_to_hz(np.linspace(hz_to_mel(f_min), hz_to_mel(f_max), n_mels + 2))
    fb = np.zeros((N_FFT // 2 + 1, n_mels))
    for m in range(n_mels):
        lo, ce, hi = f_pts[m], f_pts[m + 1], f_pts[m + 2]
        up = (freqs - lo) / (ce - lo)
        down = (hi - freqs) / (hi - ce)
        fb[:, m] = np.maximum(0.0, np.minimum(up, down)) * (2.0 / (hi - lo))
    return fb.astype(np.float32)


def log_mel(wav, sample_rate=22050, n_mels=80):
    """wav (S,) float32 -> (1 + S//256, n_mels) float32."""
    wav = np.asarray(wav, np.float32)
    x = np.pad(wav, (N_FFT // 2, N_FFT // 2), mode="reflect")
    T = 1 + len(wav) // HOP
    win = (0.5 - 0.5 * np.cos(2 * np.pi * np.arange(N_FFT) / N_FFT)).astype(np.float32)
    frames = np.stack([x[t * HOP: t * HOP + N_FFT] * win for t in range(T)])
    mag = np.abs(np.fft.rfft(frames.astype(np.float64), axis=1)).astype(np.float32)
    mel = mag @ filterbank(sample_rate, n_mels)
    return np.log(np.maximum(mel, 1e-5)).astype(np.float32)
