"""Deterministic synthetic weights, batches and dropout masks.

Everything here is a pure function of (name, seed) through numpy's PCG64, so the
golden-fixture generator (run in the build container against /root/reference),
the CPU tests, the GPU tests, ``smoke()`` and ``bench.py`` all see bit-identical
inputs without shipping 112 MB of weights.  Shapes and key names follow the
reference state dict (reference model/tacotron2.py:15-124,
model/decoder.py:11-51, model/attention.py:7-32, model/encoder.py:13-52,
model/postnet.py:5-49).
"""
from __future__ import annotations

import zlib
from dataclasses import dataclass, asdict

import numpy as np
import torch


@dataclass
class ModelArgs:
    """Constructor arguments of the reference ``Tacotron2`` (model/tacotron2.py:15-34)."""

    num_chars: int = 39
    encoded_dim: int = 512
    encoder_kernel_size: int = 5
    num_mels: int = 80
    prenet_dim: int = 256
    att_rnn_dim: int = 1024
    att_dim: int = 128
    rnn_hidden_dim: int = 1024
    postnet_dim: int = 512
    dropout: float = 0.5
    speaker_tokens: bool = False
    speaker_tokens_dim: int = 128
    num_speakers: int = 1
    controls: bool = False
    controls_dim: int = 0
    description_embeddings: bool = False
    description_embeddings_dim: int = 0

    def kwargs(self):
        return asdict(self)

    @property
    def encoded_full_dim(self) -> int:
        return self.encoded_dim + (128 if self.description_embeddings else 0)


VANILLA = ModelArgs()
# config/vanilla-lj-hifi-stop.json:46-49
HIFI = ModelArgs(speaker_tokens=True, num_speakers=4)
# config/controllable-lj-hifi-stop-speaker.json + config/descriptions-libritts.json:44-45
CONDITIONED = ModelArgs(
    speaker_tokens=True,
    num_speakers=4,
    controls=True,
    controls_dim=5,
    description_embeddings=True,
    description_embeddings_dim=768,
)


def _rng(name: str, seed: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64([zlib.crc32(name.encode()), seed]))


def param_shapes(a: ModelArgs) -> dict:
    """State-dict key -> shape, in the reference's registration order."""
    E, H, P, A, M = a.encoded_dim, a.att_rnn_dim, a.prenet_dim, a.att_dim, a.num_mels
    EF, C, R, PD, K = a.encoded_full_dim, a.controls_dim, a.rnn_hidden_dim, a.postnet_dim, a.encoder_kernel_size
    s = {}
    if a.speaker_tokens:
        s["speaker_embedding.weight"] = (a.num_speakers, E)
    s["encoder.embedding.weight"] = (a.num_chars + 1, E)
    for i in (0, 4, 8):
        s[f"encoder.convolutions.{i}.weight"] = (E, E, K)
        s[f"encoder.convolutions.{i}.bias"] = (E,)
        for f in ("weight", "bias", "running_mean", "running_var"):
            s[f"encoder.convolutions.{i + 1}.{f}"] = (E,)
    for sfx in ("", "_reverse"):
        s[f"encoder.lstm.weight_ih_l0{sfx}"] = (4 * (E // 2), E)
        s[f"encoder.lstm.weight_hh_l0{sfx}"] = (4 * (E // 2), E // 2)
        s[f"encoder.lstm.bias_ih_l0{sfx}"] = (4 * (E // 2),)
        s[f"encoder.lstm.bias_hh_l0{sfx}"] = (4 * (E // 2),)
    s["prenet.0.weight"] = (P, M)
    s["prenet.3.weight"] = (P, P)
    if a.description_embeddings:
        s["description_embeddings_linear.0.weight"] = (128, a.description_embeddings_dim)
        s["description_embeddings_linear.0.bias"] = (128,)
    s["att_encoder.weight"] = (A, EF)
    s["decoder.att_rnn.weight_ih"] = (4 * H, P + EF)
    s["decoder.att_rnn.weight_hh"] = (4 * H, H)
    s["decoder.att_rnn.bias_ih"] = (4 * H,)
    s["decoder.att_rnn.bias_hh"] = (4 * H,)
    s["decoder.attention.query_layer.weight"] = (A, H)
    s["decoder.attention.v.weight"] = (1, A)
    s["decoder.attention.location_conv.weight"] = (32, 2, 31)
    s["decoder.attention.location_dense.weight"] = (A, 32)
    s["decoder.lstm.weight_ih"] = (4 * R, H + EF + C)
    s["decoder.lstm.weight_hh"] = (4 * R, R)
    s["decoder.lstm.bias_ih"] = (4 * R,)
    s["decoder.lstm.bias_hh"] = (4 * R,)
    s["decoder.mel_out.weight"] = (M, R + EF + C)
    s["decoder.mel_out.bias"] = (M,)
    s["decoder.gate.weight"] = (1, R + EF)
    s["decoder.gate.bias"] = (1,)
    chans = [M] + [PD] * 4 + [M]
    for li, i in enumerate((0, 4, 8, 12, 16)):
        s[f"postnet.postnet.{i}.weight"] = (chans[li + 1], chans[li], 5)
        for f in ("weight", "bias", "running_mean", "running_var"):
            s[f"postnet.postnet.{i + 1}.{f}"] = (chans[li + 1],)
    return s


def synth_state_dict(a: ModelArgs, seed: int = 0, gate_shift: float = 0.0, gate_scale: float = 1.0) -> dict:
    """Seeded fp32 parameters with PyTorch-default-like scales.

    BatchNorm affine and running statistics are randomised (not the 1/0/0/1
    defaults) so an eval-mode parity test actually exercises them.
    ``gate_shift``/``gate_scale`` move ``decoder.gate`` so the stop gate fires at
    varied frames (random-init logits are ~+0.02).
    """
    sd = {}
    for name, shape in param_shapes(a).items():
        g = _rng(name, seed)
        if name.endswith("embedding.weight"):
            w = g.standard_normal(shape) * 0.5
            if name == "encoder.embedding.weight":
                w[0] = 0.0  # padding_idx=0 (model/encoder.py:25)
        elif name.endswith("running_var"):
            w = g.uniform(0.5, 1.5, shape)
        elif name.endswith("running_mean"):
            w = g.standard_normal(shape) * 0.1
        elif len(shape) == 1 and (".convolutions." in name or ".postnet." in name) and int(name.split(".")[2]) % 4 == 1:
            # BatchNorm weight / bias
            w = g.uniform(0.5, 1.5, shape) if name.endswith("weight") else g.uniform(-0.2, 0.2, shape)
        else:
            if "lstm" in name or "att_rnn" in name:
                hid = shape[0] // 4
                bound = 1.0 / np.sqrt(hid)
            elif len(shape) >= 2:
                fan_in = int(np.prod(shape[1:]))
                bound = 1.0 / np.sqrt(fan_in)
            else:
                bound = 0.05
            w = g.uniform(-bound, bound, shape)
        sd[name] = torch.from_numpy(np.ascontiguousarray(w, dtype=np.float32))
    sd["decoder.gate.weight"] = sd["decoder.gate.weight"] * gate_scale
    sd["decoder.gate.bias"] = sd["decoder.gate.bias"] * gate_scale + gate_shift
    for name in list(sd):
        if name.endswith("running_var"):
            sd[name.replace("running_var", "num_batches_tracked")] = torch.zeros((), dtype=torch.int64)
    return sd


def synth_batch(a: ModelArgs, B: int, L: int, T: int, seed: int = 1234, ragged: bool = True) -> dict:
    """Synthetic batch in the collate layout (datasets/tts_dataloader.py:8-35).

    chars_idx (B,L) int64 zero padded, mel (B,T,80) zero padded, gate (B,T,1)
    ones with the last valid frame 0 (datasets/tts_dataset.py:213-214).
    """
    g = _rng("batch", seed)
    if ragged and B > 1:
        clen = np.linspace(L, max(1, int(0.6 * L)), B).round().astype(np.int64)
        mlen = np.linspace(T, max(1, int(0.6 * T)), B).round().astype(np.int64)
        perm = g.permutation(B)  # not sorted: exercises enforce_sorted=False (model/encoder.py:61-63)
        clen, mlen = clen[perm], mlen[perm]
        clen[g.integers(0, B)] = L
        mlen[g.integers(0, B)] = T
    else:
        clen = np.full(B, L, np.int64)
        mlen = np.full(B, T, np.int64)
    chars = g.integers(1, a.num_chars + 1, (B, L)).astype(np.int64)
    chars[np.arange(L)[None, :] >= clen[:, None]] = 0
    mel = np.clip(g.standard_normal((B, T, a.num_mels)) * 2.0 - 5.0, np.log(1e-5), 2.0).astype(np.float32)
    valid = np.arange(T)[None, :] < mlen[:, None]
    mel[~valid] = 0.0
    gate = valid.astype(np.float32)
    gate[np.arange(B), mlen - 1] = 0.0
    out = {
        "chars_idx": torch.from_numpy(chars),
        "chars_idx_len": torch.from_numpy(clen),
        "mel_spectrogram": torch.from_numpy(mel),
        "mel_spectrogram_len": torch.from_numpy(mlen),
        "gate": torch.from_numpy(gate[:, :, None].copy()),
    }
    if a.speaker_tokens:
        out["speaker_id"] = torch.from_numpy(g.integers(0, a.num_speakers, (B,)).astype(np.int64))
    if a.controls:
        out["controls"] = torch.from_numpy(g.uniform(-1, 1, (B, a.controls_dim)).astype(np.float32))
    if a.description_embeddings:
        out["description_embeddings"] = torch.from_numpy(
            np.tanh(g.standard_normal((B, a.description_embeddings_dim))).astype(np.float32)
        )
    return out


def synth_masks(a: ModelArgs, B: int, L: int, T: int, seed: int = 7, train: bool = False) -> dict:
    """Keep-masks (uint8, 1 = keep) in the layout the C-ABI side channel takes.

    prenet: (T+1, 2, B, P) -- step t uses [t]; the reference draws these from the
    global RNG (model/modules.py:10-12), here they are injected.  Train-mode only:
    att_h / rnn_h (T, B, H) p=0.1 (model/decoder.py:29,43), encoder (3, B, L, E),
    postnet (5, B, T, 512) (last layer uses the first num_mels channels).
    """
    g = _rng("masks", seed)
    m = {"prenet": torch.from_numpy((g.random((T + 1, 2, B, a.prenet_dim)) >= a.dropout).astype(np.uint8))}
    if train:
        m["att_h"] = torch.from_numpy((g.random((T, B, a.att_rnn_dim)) >= 0.1).astype(np.uint8))
        m["rnn_h"] = torch.from_numpy((g.random((T, B, a.rnn_hidden_dim)) >= 0.1).astype(np.uint8))
        m["encoder"] = torch.from_numpy((g.random((3, B, L, a.encoded_dim)) >= a.dropout).astype(np.uint8))
        m["postnet"] = torch.from_numpy((g.random((5, B, T, a.postnet_dim)) >= a.dropout).astype(np.uint8))
    return m


MEL_GOLDEN_LENGTHS = (22050, 5000, 1024, 777, 256 * 40)  # samples; tests/golden/melspec.npz


def synth_wav(n_samples: int, seed: int = 0) -> np.ndarray:
    """Speech-like synthetic waveform: a few decaying harmonics with vibrato + band noise + a silent tail, in [-1, 1]."""
    g = _rng("wav", seed)
    t = np.arange(n_samples) / 22050.0
    f0 = 110.0 + 40.0 * g.random() + 8.0 * np.sin(2 * np.pi * 5.0 * t)
    ph = 2 * np.pi * np.cumsum(f0) / 22050.0
    x = sum((0.6 ** h) * np.sin((h + 1) * ph + g.random() * 6.28) for h in range(12))
    env = 0.5 + 0.5 * np.sin(2 * np.pi * 2.3 * t + g.random() * 6.28)
    x = 0.2 * x * env + 0.02 * g.standard_normal(n_samples)
    x[int(0.9 * n_samples):] *= 0.001  # near-silence: exercises the log floor
    return np.clip(x, -1, 1).astype(np.float32)
