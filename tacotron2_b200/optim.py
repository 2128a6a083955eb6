"""Optimizer shell of the training step: global-norm gradient clipping + Adam with L2 weight decay.

Semantics of ``torch.nn.utils.clip_grad_norm_(params, max_norm)`` followed by ``torch.optim.Adam(params, lr, betas, eps,
weight_decay).step()`` -- what the reference configures in ``model/tts_model.py:78-91`` (Adam, lr 1e-3, weight_decay 1e-6) and
``run/train.py:239-240`` (``gradient_clip_val=1.0``) -- as three CUDA launches (``t2_clip_adam_step``) with no host
synchronisation.  ``MultiStepLR`` (``model/tts_model.py:85-90``) is host arithmetic: :meth:`ClipAdam.set_lr`.
"""
from __future__ import annotations

import ctypes as C

import torch
from torch.autograd.graph import increment_version

from . import _lib


class ClipAdam:
    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0, max_norm: float = 1.0):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("ClipAdam: no trainable parameters")
        for p in self.params:
            if p.device.type != "cuda" or p.dtype != torch.float32 or not p.is_contiguous():
                raise RuntimeError("tacotron2_b200 runs on contiguous fp32 CUDA tensors only; there is no CPU fallback")
        self.lr, self.betas, self.eps, self.weight_decay, self.max_norm = float(lr), (float(betas[0]), float(betas[1])), float(eps), \
            float(weight_decay), float(max_norm)
        self.step_count = 0
        dev = self.params[0].device
        self.exp_avg = [torch.zeros_like(p) for p in self.params]
        self.exp_avg_sq = [torch.zeros_like(p) for p in self.params]
        self.grad_norm = torch.zeros((), device=dev)
        n = len(self.params)
        self._numel = (C.c_longlong * n)(*[p.numel() for p in self.params])
        lib = _lib.load()
        chunks = lib.t2_clip_adam_chunks(n, self._numel)
        self._ws = torch.empty(lib.t2_clip_adam_workspace_bytes(n, chunks), dtype=torch.uint8, device=dev)
        self._table_cached = 0

    def set_lr(self, lr: float) -> None:
        self.lr = float(lr)

    def state_dict(self) -> dict:
        """The layout of ``torch.optim.Adam.state_dict()`` (what a Lightning checkpoint stores under ``optimizer_states``), so a
        run can be handed back to the reference's Trainer: state[i] = {step, exp_avg, exp_avg_sq}, one param group."""
        n = len(self.params)
        state = {i: {"step": torch.tensor(float(self.step_count)), "exp_avg": self.exp_avg[i], "exp_avg_sq": self.exp_avg_sq[i]}
                 for i in range(n)} if self.step_count > 0 else {}
        group = {"lr": self.lr, "betas": self.betas, "eps": self.eps, "weight_decay": self.weight_decay, "amsgrad": False,
                 "maximize": False, "foreach": None, "capturable": False, "differentiable": False, "fused": None,
                 "decoupled_weight_decay": False, "params": list(range(n))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd: dict) -> None:
        group = sd["param_groups"][0]
        if len(group["params"]) != len(self.params):
            raise ValueError(f"ClipAdam.load_state_dict: {len(group['params'])} parameters in the checkpoint, {len(self.params)} here")
        self.lr, self.betas, self.eps = float(group["lr"]), tuple(float(b) for b in group["betas"]), float(group["eps"])
        self.weight_decay = float(group["weight_decay"])
        steps = set()
        for j, idx in enumerate(group["params"]):
            st = sd["state"].get(idx)
            if st is None:
                self.exp_avg[j].zero_(), self.exp_avg_sq[j].zero_()
                continue
            self.exp_avg[j].copy_(st["exp_avg"])
            self.exp_avg_sq[j].copy_(st["exp_avg_sq"])
            steps.add(int(float(st["step"])))
        if len(steps) > 1:
            raise ValueError("ClipAdam.load_state_dict: per-parameter step counts differ")
        self.step_count = steps.pop() if steps else 0

    def zero_grad(self, set_to_none: bool = True) -> None:
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self) -> torch.Tensor:
        """One update; returns the (device) pre-clip global gradient norm.  Parameters without a gradient take a zero gradient
        (torch skips them; with weight_decay == 0 and zero moments the two coincide, and the reference always has all gradients)."""
        lib = _lib.load()
        n = len(self.params)
        grads = []
        for p in self.params:
            g = p.grad
            if g is None:
                g = torch.zeros_like(p)
            elif not g.is_contiguous() or g.dtype != torch.float32:
                g = g.contiguous().float()
            grads.append(g)
        arr = lambda ts: (C.c_void_p * n)(*[t.data_ptr() for t in ts])  # noqa: E731
        self.step_count += 1
        dev = self.params[0].device
        with torch.cuda.device(dev):
            st = torch.cuda.current_stream().cuda_stream
            _lib.check(lib.t2_clip_adam_step(n, arr(self.params), arr(grads), arr(self.exp_avg), arr(self.exp_avg_sq), self._numel, self.lr,
                                             self.betas[0], self.betas[1], self.eps, self.weight_decay, self.max_norm, self.step_count,
                                             self._table_cached, _lib.ptr(self.grad_norm), _lib.ptr(self._ws), self._ws.numel(), st),
                       "t2_clip_adam_step")
        self._table_cached = 1
        # the kernels wrote the parameters behind torch's back: bump the version counters that Tacotron2._ready watches to
        # re-pack the bf16 (hi,lo) operand copies (no device work)
        increment_version(self.params)
        return self.grad_norm
