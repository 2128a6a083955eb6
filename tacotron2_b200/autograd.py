"""torch.autograd glue for the training path.

Each ``Function`` below is a thin binding: forward calls a ``t2_*_train_fwd`` entry point of libt2b200 that records a
tape, backward calls the matching ``t2_*_bwd``.  No arithmetic happens in Python.  Parameters are passed to ``apply`` so
autograd routes their gradients; the C side writes them into a freshly allocated flat buffer per module ("bucket"),
registered key by key with ``t2_ctx_set_grad``.  A bucket is also the unit of the data-parallel gradient all-reduce
(``tacotron2_b200.parallel``): its NCCL all-reduce is enqueued the moment the module's backward has been launched, so it
overlaps with the backward of the modules upstream.
"""
from __future__ import annotations

import torch

from . import _lib

P = _lib.ptr


def _stream():
    import ctypes as C
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _bucket(owner, lib, params: dict):
    """Allocate one flat gradient buffer for ``params`` (key -> Parameter), register the slices, return views in order."""
    n = sum(p.numel() for p in params.values())
    dev = next(iter(params.values())).device
    flat = torch.empty(n, dtype=torch.float32, device=dev)
    views, off = [], 0
    for k, p in params.items():
        v = flat[off:off + p.numel()]
        _lib.check(lib.t2_ctx_set_grad(owner._ctx, k.encode(), P(v), p.numel()), "t2_ctx_set_grad")
        views.append(v.view_as(p))
        off += p.numel()
    return flat, views


def _take_tape(owner, nbytes: int, dev):
    """The decoder tape is several GB per 64-utterance chunk: keep the buffers on the model between steps instead of going
    through the allocator every step.  A buffer is checked out by a forward and handed back by its backward."""
    pool = owner.__dict__.setdefault("_tape_pool", [])
    for i, t in enumerate(pool):
        if t.device == dev and t.numel() >= nbytes:
            return pool.pop(i)
    return torch.empty(nbytes, dtype=torch.uint8, device=dev)


def _give_tape(owner, tape):
    pool = owner.__dict__.setdefault("_tape_pool", [])
    if len(pool) < 8:
        pool.append(tape)


def _count_bn_batches(module):
    """nn.BatchNorm1d increments num_batches_tracked on every train-mode forward; the kernels update the running statistics in
    place, the counter is bumped here so that saved checkpoints match the reference's."""
    for m in module.modules():
        if isinstance(m, torch.nn.BatchNorm1d) and m.num_batches_tracked is not None:
            m.num_batches_tracked += 1


def _after_bucket(owner, flat, group, views):
    hook = getattr(owner, "_grad_bucket_hook", None)
    if hook is not None:  # data-parallel: async all-reduce of this bucket (parallel.py)
        hook(flat, list(owner._group_params(group).values()), views)


class PostnetFn(torch.autograd.Function):
    """Postnet.forward under train() (model/postnet.py:8-52) on channels-last (B,T,M)."""

    @staticmethod
    def forward(ctx, owner, keep, seed, mels_cl, *params):
        lib = owner._ready(mels_cl.device)
        B, T, M = mels_cl.shape
        x = mels_cl.contiguous()
        tape = torch.empty(lib.t2_postnet_tape_bytes(owner._ctx, B, T), dtype=torch.uint8, device=x.device)
        ws = owner._train_workspace(lib, x.device, B, 1, T)
        out = torch.empty(B, T, M, device=x.device)
        _lib.check(lib.t2_postnet_train_fwd(owner._ctx, P(x), T, B, T, P(keep), seed, P(out), P(tape), tape.numel(), P(ws), ws.numel(),
                                            _stream()), "t2_postnet_train_fwd")
        owner._bn_dirty = True  # running statistics changed in place: re-fold them before the next eval() forward
        _count_bn_batches(owner.postnet)
        ctx.owner, ctx.keep, ctx.seed, ctx.tape, ctx.shape = owner, keep, seed, tape, (B, T, M)
        return out

    @staticmethod
    def backward(ctx, d_out):
        owner = ctx.owner
        B, T, M = ctx.shape
        lib = owner._ready(d_out.device)
        flat, views = _bucket(owner, lib, owner._group_params("postnet"))
        ws = owner._train_workspace(lib, d_out.device, B, 1, T)
        d_mels = torch.empty(B, T, M, device=d_out.device)
        _lib.check(lib.t2_postnet_bwd(owner._ctx, B, T, P(ctx.keep), ctx.seed, P(d_out.contiguous()), None, P(ctx.tape), ctx.tape.numel(),
                                      P(d_mels), P(ws), ws.numel(), _stream()), "t2_postnet_bwd")
        _after_bucket(owner, flat, "postnet", views)
        return (None, None, None, d_mels, *views)


class EncoderFn(torch.autograd.Function):
    """Encoder.forward under train() (model/encoder.py:54-67)."""

    @staticmethod
    def forward(ctx, owner, chars, lens, keep, seed, *params):
        dev = chars.device
        lib = owner._ready(dev)
        B, L = chars.shape
        tape = torch.empty(lib.t2_encoder_tape_bytes(owner._ctx, B, L), dtype=torch.uint8, device=dev)
        ws = owner._train_workspace(lib, dev, B, L, 1)
        out = torch.empty(B, L, owner.embedding_dim, device=dev)
        _lib.check(lib.t2_encoder_train_fwd(owner._ctx, P(chars), P(lens), B, L, P(keep), seed, P(out), P(tape), tape.numel(), P(ws),
                                            ws.numel(), _stream()), "t2_encoder_train_fwd")
        owner._bn_dirty = True
        _count_bn_batches(owner.encoder)
        ctx.owner, ctx.args, ctx.tape, ctx.out = owner, (chars, lens, keep, seed, B, L), tape, out
        return out

    @staticmethod
    def backward(ctx, d_out):
        owner = ctx.owner
        chars, lens, keep, seed, B, L = ctx.args
        lib = owner._ready(chars.device)
        flat, views = _bucket(owner, lib, owner._group_params("encoder"))
        ws = owner._train_workspace(lib, chars.device, B, L, 1)
        _lib.check(lib.t2_encoder_bwd(owner._ctx, P(chars), P(lens), B, L, P(keep), seed, P(ctx.out), P(d_out.contiguous()), P(ctx.tape),
                                      ctx.tape.numel(), P(ws), ws.numel(), _stream()), "t2_encoder_bwd")
        _after_bucket(owner, flat, "encoder", views)
        return (None,) * 5 + tuple(views)


class MemoryFn(torch.autograd.Function):
    """Conditioning + att_encoder (model/tacotron2.py:201-212,229): encoded -> (memory, processed_memory)."""

    @staticmethod
    def forward(ctx, owner, spk, desc, encoded, *params):
        memory, pmem = owner._memory(encoded, spk, desc)
        ctx.owner, ctx.args = owner, (spk, desc, memory)
        return memory, pmem

    @staticmethod
    def backward(ctx, d_memory, d_pmem):
        owner = ctx.owner
        spk, desc, memory = ctx.args
        dev = memory.device
        B, L, _ = memory.shape
        lib = owner._ready(dev)
        flat, views = _bucket(owner, lib, owner._group_params("memory"))
        ws = owner._train_workspace(lib, dev, B, L, 1)
        z = lambda t, like: torch.zeros_like(like) if t is None else t.contiguous()  # noqa: E731
        d_memory = z(d_memory, memory)
        d_pmem = torch.zeros(B, L, owner.att_dim, device=dev) if d_pmem is None else d_pmem.contiguous()
        d_enc = torch.empty(B, L, owner.embedding_dim, device=dev)
        spk_c = None if spk is None else spk.to(device=dev, dtype=torch.int64).contiguous()
        desc_c = None if desc is None else desc.to(device=dev, dtype=torch.float32).contiguous()
        _lib.check(lib.t2_memory_bwd(owner._ctx, P(memory), P(spk_c), P(desc_c), P(d_memory), P(d_pmem), B, L, P(d_enc), P(ws), ws.numel(),
                                     _stream()), "t2_memory_bwd")
        _after_bucket(owner, flat, "memory", views)
        return (None, None, None, d_enc, *views)


class DecodeTfFn(torch.autograd.Function):
    """The teacher-forced frame loop under train() (model/tacotron2.py:255-258,276-317) for <= 64 utterances."""

    @staticmethod
    def forward(ctx, owner, mask, ctrl, target, pk, ak, rk, seed, memory, pmem, *params):
        dev = memory.device
        lib = owner._ready(dev)
        B, L, _ = memory.shape
        T, M = target.shape[1], target.shape[2]
        memory, pmem = memory.contiguous(), pmem.contiguous()
        tape = _take_tape(owner, lib.t2_decode_tape_bytes(owner._ctx, B, L, T), dev)
        ws = owner._train_workspace(lib, dev, B, L, T)
        mels, gates, aligns = torch.empty(B, T, M, device=dev), torch.empty(B, T, device=dev), torch.empty(B, T, L, device=dev)
        _lib.check(lib.t2_decode_tf_train_fwd(owner._ctx, P(memory), P(pmem), P(mask), P(ctrl), P(target), B, L, T, P(pk), P(ak), P(rk), seed,
                                              P(mels), P(gates), P(aligns), P(tape), tape.numel(), P(ws), ws.numel(), _stream()),
                   "t2_decode_tf_train_fwd")
        ctx.owner, ctx.tape, ctx.seed, ctx.dims = owner, tape, seed, (B, L, T, M)
        ctx.keep = (mask, target, ak, rk, memory, pmem, aligns)
        ctx.mark_non_differentiable(aligns)
        return mels, gates, aligns

    @staticmethod
    def backward(ctx, d_mels, d_gates, _d_aligns):
        owner = ctx.owner
        B, L, T, M = ctx.dims
        mask, target, ak, rk, memory, pmem, aligns = ctx.keep
        dev = memory.device
        lib = owner._ready(dev)
        flat, views = _bucket(owner, lib, owner._group_params("decoder"))
        ws = owner._train_workspace(lib, dev, B, L, T)
        z = lambda t, *s: torch.zeros(*s, device=dev) if t is None else t.contiguous()  # noqa: E731
        d_mels, d_gates = z(d_mels, B, T, M), z(d_gates, B, T)
        d_memory, d_pmem = torch.empty_like(memory), torch.empty_like(pmem)
        _lib.check(lib.t2_decode_tf_bwd(owner._ctx, P(memory), P(pmem), P(mask), P(target), B, L, T, P(ak), P(rk), ctx.seed, P(aligns),
                                        P(d_mels), P(d_gates), P(ctx.tape), ctx.tape.numel(), P(d_memory), P(d_pmem), P(ws), ws.numel(),
                                        _stream()), "t2_decode_tf_bwd")
        _after_bucket(owner, flat, "decoder", views)
        _give_tape(owner, ctx.tape)   # stream-ordered: the next forward's writes are enqueued behind this backward
        ctx.tape = None
        return (None,) * 8 + (d_memory, d_pmem, *views)


class FinalizeFn(torch.autograd.Function):
    """mels_post = mels + postnet_out, then the three output masks (model/tacotron2.py:333-345)."""

    @staticmethod
    def forward(ctx, owner, lengths, mels_raw, post_out, gates_raw):
        lib = owner._ready(mels_raw.device)
        B, T, M = mels_raw.shape
        mels, post, gates = mels_raw.clone(), post_out.clone(), gates_raw.clone()
        _lib.check(lib.t2_finalize(owner._ctx, P(mels), P(post), P(gates), P(lengths), B, T, T, _stream()), "t2_finalize")
        ctx.owner, ctx.lengths, ctx.shape = owner, lengths, (B, T, M)
        return mels, post, gates

    @staticmethod
    def backward(ctx, d_mels, d_post, d_gates):
        owner = ctx.owner
        B, T, M = ctx.shape
        lib = _lib.load()
        dev = ctx.lengths.device
        c = lambda t: None if t is None else t.contiguous()  # noqa: E731
        d_raw, d_po, d_g = torch.empty(B, T, M, device=dev), torch.empty(B, T, M, device=dev), torch.empty(B, T, device=dev)
        _lib.check(lib.t2_finalize_bwd(owner._ctx, P(c(d_mels)), P(c(d_post)), P(c(d_gates)), P(ctx.lengths), B, T, P(d_raw), P(d_po),
                                       P(d_g), _stream()), "t2_finalize_bwd")
        return None, None, d_raw, d_po, d_g


class LossFn(torch.autograd.Function):
    """training_step loss (model/tts_model.py:197-201) with its gradient computed in the same kernel."""

    @staticmethod
    def forward(ctx, owner, mels, post, gates, mel_target, gate_target):
        lib = owner._ready(mels.device)
        B, T, M = mels.shape
        dev = mels.device
        loss = torch.empty((), device=dev)
        d_m, d_p, d_g = torch.empty(B, T, M, device=dev), torch.empty(B, T, M, device=dev), torch.empty(B, T, 1, device=dev)
        ws = owner._train_workspace(lib, dev, B, 1, T)
        f = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
        ins = [f(mels), f(post), f(gates), f(mel_target), f(gate_target)]   # alive until after the launch (no freed temporaries)
        _lib.check(lib.t2_loss_fwd_bwd(owner._ctx, *[P(t) for t in ins], B, T, P(loss),
                                       P(d_m), P(d_p), P(d_g), P(ws), ws.numel(), _stream()), "t2_loss_fwd_bwd")
        del ins
        ctx.save_for_backward(d_m, d_p, d_g)
        return loss

    @staticmethod
    def backward(ctx, g):
        d_m, d_p, d_g = ctx.saved_tensors
        return None, d_m * g, d_p * g, d_g * g, None, None
