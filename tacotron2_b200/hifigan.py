"""HiFi-GAN generator behind the reference's module API, backed by libt2b200.

``Generator(h)`` keeps the constructor argument (the attribute-dict HiFi-GAN config of ``run/say.py:76-84``), the ``forward(x)``
signature ``(B, 80, T) -> (B, 1, T * prod(upsample_rates))``, ``remove_weight_norm()`` and the parameter names of
``model/hifi_gan.py:154-216`` (``conv_pre``, ``ups.i``, ``resblocks.n.convs1.m`` / ``convs2.m``, ``conv_post``), so the reference's
call site ``generator(mel_spectrogram_post[:, :-1].swapaxes(1, 2))`` (``run/say.py:153-157``) works unchanged and a reference
generator checkpoint loads with ``load_state_dict`` -- with or without weight normalisation: ``weight_g`` / ``weight_v`` pairs are
folded into ``weight`` on load (what ``remove_weight_norm`` computes).  The ``torch.nn`` sub-modules are parameter containers; all
arithmetic runs in ``t2_hifigan_fwd`` (tcgen05 implicit-GEMM convs, ``csrc/hifigan.cu``).  CUDA tensors only; no fallback.
"""
from __future__ import annotations

import ctypes as C
import os

import torch
from torch import nn

from . import _lib

LRELU_SLOPE = 0.1


def _get(h, k):
    return h[k] if isinstance(h, dict) else getattr(h, k)


def get_padding(kernel_size, dilation=1):
    return int((kernel_size * dilation - dilation) / 2)


class ResBlock1(nn.Module):
    """model/hifi_gan.py:20-112 (parameter container)."""

    def __init__(self, h, channels, kernel_size=3, dilation=(1, 3, 5)):
        super().__init__()
        self.convs1 = nn.ModuleList([nn.Conv1d(channels, channels, kernel_size, 1, dilation=d, padding=get_padding(kernel_size, d)) for d in dilation])
        self.convs2 = nn.ModuleList([nn.Conv1d(channels, channels, kernel_size, 1, dilation=1, padding=get_padding(kernel_size, 1)) for _ in dilation])
        for m in list(self.convs1) + list(self.convs2):
            m.weight.data.normal_(0.0, 0.01)


class ResBlock2(nn.Module):
    """model/hifi_gan.py:115-151 (parameter container)."""

    def __init__(self, h, channels, kernel_size=3, dilation=(1, 3)):
        super().__init__()
        self.convs = nn.ModuleList([nn.Conv1d(channels, channels, kernel_size, 1, dilation=d, padding=get_padding(kernel_size, d)) for d in dilation])
        for m in self.convs:
            m.weight.data.normal_(0.0, 0.01)


class Generator(nn.Module):
    def __init__(self, h):
        super().__init__()
        self.h = h
        rk, ur, uk = list(_get(h, "resblock_kernel_sizes")), list(_get(h, "upsample_rates")), list(_get(h, "upsample_kernel_sizes"))
        rd = [list(d) for d in _get(h, "resblock_dilation_sizes")]
        c0 = int(_get(h, "upsample_initial_channel"))
        self.num_kernels, self.num_upsamples = len(rk), len(ur)
        self.resblock_kind = 1 if str(_get(h, "resblock")) == "1" else 2
        self.conv_pre = nn.Conv1d(80, c0, 7, 1, padding=3)
        block = ResBlock1 if self.resblock_kind == 1 else ResBlock2
        self.ups = nn.ModuleList([nn.ConvTranspose1d(c0 // (2 ** i), c0 // (2 ** (i + 1)), k, u, padding=(k - u) // 2) for i, (u, k) in enumerate(zip(ur, uk))])
        self.resblocks = nn.ModuleList()
        ch = c0
        for i in range(len(self.ups)):
            ch = c0 // (2 ** (i + 1))
            for k, d in zip(rk, rd):
                self.resblocks.append(block(h, ch, k, d))
        self.conv_post = nn.Conv1d(ch, 1, 7, 1, padding=3)
        for m in list(self.ups) + [self.conv_post]:
            m.weight.data.normal_(0.0, 0.01)
        cfg = _lib.T2HifiganConfig()
        cfg.num_mels, cfg.upsample_initial_channel, cfg.num_upsamples = 80, c0, len(ur)
        cfg.resblock, cfg.num_kernels = self.resblock_kind, len(rk)
        for i, (u, k) in enumerate(zip(ur, uk)):
            cfg.upsample_rates[i], cfg.upsample_kernel_sizes[i] = int(u), int(k)
        nd = 3 if self.resblock_kind == 1 else 2
        for j, (k, d) in enumerate(zip(rk, rd)):
            assert len(d) == nd, f"ResBlock{self.resblock_kind} takes {nd} dilations per kernel size"
            cfg.resblock_kernel_sizes[j] = int(k)
            for m in range(nd):
                cfg.resblock_dilation_sizes[j][m] = int(d[m])
        cfg.precision = 1 if os.environ.get("T2B200_PRECISION", "fp32").lower() == "bf16" else 0
        self._cfg, self._ctx, self._sig, self._ws = cfg, None, None, None
        self.upsample_factor = 1
        for u in ur:
            self.upsample_factor *= int(u)

    # ------------------------------------------------------------------ reference API
    def remove_weight_norm(self):
        """run/say.py:88.  Weight normalisation is folded at load time (see ``load_state_dict``); nothing to remove."""
        return self

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        """Accepts the reference generator's checkpoint in either form: plain ``weight`` (after ``remove_weight_norm``) or the
        ``weight_g`` / ``weight_v`` pair of ``torch.nn.utils.weight_norm`` (dim 0), folded as w = v * g / ||v||."""
        sd = dict(state_dict)
        for k in [k for k in sd if k.endswith(".weight_v")]:
            base = k[: -len("_v")]
            v, g = sd.pop(k), sd.pop(base + "_g")
            norm = v.flatten(1).norm(dim=1).reshape(-1, *([1] * (v.dim() - 1)))
            sd[base] = v * (g / norm)
        return super().load_state_dict(sd, strict=strict, assign=assign)

    def __del__(self):
        try:
            if self._ctx is not None:
                _lib.load().t2_hifigan_destroy(self._ctx)
        except Exception:
            pass

    def _ready(self, device):
        if device.type != "cuda":
            raise RuntimeError("tacotron2_b200.hifigan runs on CUDA (sm_100a) tensors only; there is no CPU fallback")
        lib = _lib.load()
        tensors = dict(self.named_parameters())
        for k, v in tensors.items():
            if v.device != device or v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be a contiguous fp32 tensor on {device}")
        sig = tuple((k, v.data_ptr(), v._version) for k, v in tensors.items())
        with torch.cuda.device(device):
            if self._ctx is None:
                hdl = C.c_void_p()
                _lib.check(lib.t2_hifigan_create(C.byref(self._cfg), C.byref(hdl)), "t2_hifigan_create")
                self._ctx = hdl
            if sig != self._sig:
                for k, v in tensors.items():
                    _lib.check(lib.t2_hifigan_set_param(self._ctx, k.encode(), _lib.ptr(v), v.numel()), "t2_hifigan_set_param")
                _lib.check(lib.t2_hifigan_commit(self._ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "t2_hifigan_commit")
                self._sig = sig
        return lib

    def forward(self, x):
        """x (B, 80, T) -> (B, 1, T * prod(upsample_rates)), model/hifi_gan.py:192-209."""
        dev = x.device
        lib = self._ready(dev)
        xin = x.to(torch.float32).contiguous()
        B, M, T = xin.shape
        assert M == 80, "conv_pre takes 80 mel channels (model/hifi_gan.py:161)"
        need = lib.t2_hifigan_workspace_bytes(self._ctx, B, T)
        if need < 0:
            raise RuntimeError("t2_hifigan_workspace_bytes failed")
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        wav = torch.empty(B, 1, T * self.upsample_factor, device=dev)
        _lib.check(lib.t2_hifigan_fwd(self._ctx, _lib.ptr(xin), B, T, _lib.ptr(wav), _lib.ptr(self._ws), self._ws.numel(),
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream)), "t2_hifigan_fwd")
        return wav


UNIVERSAL_V1 = {   # the config run/say.py:76 loads (web_checkpoints/hifi-gan/UNIVERSAL_V1/config.json): HiFi-GAN V1 generator
    "resblock": "1", "upsample_rates": [8, 8, 2, 2], "upsample_kernel_sizes": [16, 16, 4, 4], "upsample_initial_channel": 512,
    "resblock_kernel_sizes": [3, 7, 11], "resblock_dilation_sizes": [[1, 3, 5], [1, 3, 5], [1, 3, 5]],
}


def generator_flops_per_frame(h) -> float:
    """Multiply-add FLOPs (x2) of Generator.forward per input mel frame."""
    rk, ur, uk = list(_get(h, "resblock_kernel_sizes")), list(_get(h, "upsample_rates")), list(_get(h, "upsample_kernel_sizes"))
    nd = 3 if str(_get(h, "resblock")) == "1" else 2
    convs_per_dil = 2 if str(_get(h, "resblock")) == "1" else 1
    c = int(_get(h, "upsample_initial_channel"))
    fl, rate = 2.0 * 80 * c * 7, 1
    for u, k in zip(ur, uk):
        fl += 2.0 * c * (c // 2) * k * rate          # transposed conv: k / u taps per output position, u * rate positions per frame
        rate *= u
        c //= 2
        fl += sum(2.0 * c * c * kk * nd * convs_per_dil for kk in rk) * rate
    return fl + 2.0 * c * 7 * rate


def synth_generator_state(h, seed: int = 0, scale: float = 0.8):
    """Seeded weights in the reference's (weight-norm-free) key layout, at a scale that keeps activations O(1) through the stack."""
    import numpy as np
    g = Generator(h)
    rng = np.random.Generator(np.random.PCG64(seed))
    sd = {}
    for k, v in g.state_dict().items():
        if k.endswith("weight"):
            fan_in = v.shape[1] * v.shape[2] if not k.startswith("ups.") else v.shape[0] * v.shape[2] / _up_rate(h, k)
            sd[k] = torch.from_numpy((rng.standard_normal(tuple(v.shape)) * scale / np.sqrt(fan_in)).astype(np.float32))
        else:
            sd[k] = torch.from_numpy((rng.standard_normal(tuple(v.shape)) * 0.05).astype(np.float32))
    return sd


def _up_rate(h, key):
    return list(_get(h, "upsample_rates"))[int(key.split(".")[1])]


def bench(dev, B: int = 8, T: int = 800, steps: int = 3):
    """Generator.forward on B x T mel frames (vocoding one C3-sized slice), resident inputs: ms, frames/s, achieved TFLOP/s against the
    measured dense bf16 peak (fp32-parity arithmetic issues three bf16 passes: both rates are reported)."""
    import json
    g = Generator(UNIVERSAL_V1).to(dev).eval()
    g.load_state_dict(synth_generator_state(UNIVERSAL_V1, seed=0))
    x = torch.randn(B, 80, T, device=dev)
    with torch.no_grad():
        for _ in range(2):
            y = g(x)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            y = g(x)
        e1.record()
        e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    assert y.shape == (B, 1, T * 256) and bool(torch.isfinite(y).all())
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        pk = json.load(open(os.path.join(root, "MEASURED_PEAKS.json")))
        peak = float(pk.get("bf16_tflops_sustained", pk.get("bf16_tflops", 1342.0)))
    except Exception:  # noqa: BLE001
        peak = 1342.0
    tf = generator_flops_per_frame(UNIVERSAL_V1) * B * T / (ms * 1e-3) / 1e12
    return {"workload": f"HiFi-GAN V1 generator (UNIVERSAL_V1 config), {B} x {T} mel frames -> {B} x {T * 256} samples, fp32-parity = 3 bf16 passes",
            "ms": ms, "mel_frames_per_s": B * T / (ms * 1e-3), "samples_per_s": B * T * 256 / (ms * 1e-3),
            "mflop_per_frame": generator_flops_per_frame(UNIVERSAL_V1) / 1e6, "algorithmic_tflops": tf, "issued_bf16_tflops": 3 * tf,
            "bound": "tensor", "peak_tflops": peak, "frac_algorithmic": tf / peak, "frac_issued": 3 * tf / peak}
