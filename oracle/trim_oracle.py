"""CPU restatement of ``librosa.effects.trim`` as the reference calls it (datasets/tts_dataset.py:194-200).  TEST INFRASTRUCTURE ONLY.

librosa is an external, un-vendored dependency of the reference and is absent from this image: **parity unpinned vs librosa**.
The algorithm restated here is librosa's published one (librosa/effects.py ``trim`` / ``_signal_to_frame_nonsilent``, 0.10):
``rms = feature.rms(y, frame_length, hop_length)`` on centred frames with zero padding, ``db = amplitude_to_db(rms, ref=np.max,
top_db=None)`` with amin 1e-5, non-silent = ``db > -top_db``; the kept interval runs from the first non-silent frame's first
sample to ``min(len, (last + 1) * hop)``.  Returns (trimmed signal, (start, end))."""
import numpy as np


def trim(y: np.ndarray, top_db: float = 60.0, frame_length: int = 2048, hop_length: int = 512):
    y = np.asarray(y, dtype=np.float32)
    n = y.shape[-1]
    pad = frame_length // 2
    yp = np.pad(y, (pad, pad), mode="constant")
    nf = 1 + n // hop_length
    idx = np.arange(frame_length)[None, :] + hop_length * np.arange(nf)[:, None]
    frames = yp[np.minimum(idx, yp.shape[0] - 1)] * (idx < yp.shape[0])
    rms = np.sqrt(np.mean(frames.astype(np.float32) ** 2, axis=1, dtype=np.float32))
    amin = np.float32(1e-5)
    db = 20.0 * np.log10(np.maximum(amin, rms)) - 20.0 * np.log10(np.maximum(amin, rms.max() if nf else amin))
    nz = np.flatnonzero(db > -top_db)
    if nz.size == 0:
        return y[0:0], (0, 0)
    start, end = int(nz[0]) * hop_length, min(n, (int(nz[-1]) + 1) * hop_length)
    return y[start:end], (start, end)
