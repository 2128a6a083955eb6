"""Golden log-mels from torchaudio (build container only): the exact transform the reference states at
datasets/prosody_dataset.py:39-50,66-67.  Waveforms are regenerated from seeds by the tests (synth_wav)."""
import os
import sys

import numpy as np
import torch
import torchaudio

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
from tacotron2_b200 import synth  # noqa: E402

if __name__ == "__main__":
    tr = torchaudio.transforms.MelSpectrogram(sample_rate=22050, n_fft=1024, win_length=1024, hop_length=256, f_min=0, f_max=8000,
                                              n_mels=80, power=1, mel_scale="slaney", norm="slaney")
    out = {"fb": tr.mel_scale.fb.numpy()}
    for i, S in enumerate(synth.MEL_GOLDEN_LENGTHS):
        wav = torch.from_numpy(synth.synth_wav(S, seed=i))
        out[f"mel{i}"] = torch.log(torch.clamp(tr(wav), min=1e-5)).T.numpy()
    np.savez_compressed(os.path.join(HERE, "melspec.npz"), **out)
    print({k: v.shape for k, v in out.items()})
