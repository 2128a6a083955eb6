"""Host-side mirror of the reference module API, backed by libt2b200 (hand-written sm_100a CUDA).

Same class names, constructor arguments, forward signatures, return tuples and
state-dict keys as mattm458/tacotron2 (``model/tacotron2.py:14-347``, ``model/encoder.py``,
``model/decoder.py``, ``model/attention.py``, ``model/postnet.py``, ``model/modules.py``), so a
reference checkpoint loads with ``load_state_dict`` and the reference's callers
(``model/tts_model.py:105``, ``run/say.py:141``) can use this class unchanged.

The ``torch.nn`` sub-modules below are PARAMETER CONTAINERS only (they give the reference's
parameter names/shapes/initialisation); their own ``forward`` is never used.  All arithmetic
goes through the C ABI in ``include/t2b200.h``.  CUDA tensors only; no fallback.

Extension over the reference signature: ``dropout_masks`` (dict of uint8 keep-masks in the
layouts of ``tacotron2_b200.synth.synth_masks``) injects the dropout draws for parity tests;
when absent the kernels draw their own (counter-based RNG seeded from torch's generator).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
from torch import Tensor, nn

from . import _lib

_MAX_DECODE_BATCH = 64  # DEC_BP in csrc/decoder.cuh


class AlwaysDropout(nn.Dropout):
    """model/modules.py:4-12 -- dropout that stays active in eval().  Parameter-free; inside
    Tacotron2 its arithmetic is fused into the prenet kernels."""

    def __init__(self, p):
        super().__init__(p)

    def forward(self, X):
        self.training = True
        return super().forward(X)


def _owner_of(mod):
    o = getattr(mod, "_t2_owner", None)
    o = o() if o is not None else None
    if o is None:
        raise RuntimeError(
            f"{type(mod).__name__}.forward needs the owning tacotron2_b200.Tacotron2 (its CUDA context holds the packed "
            "weights); construct the sub-module through Tacotron2"
        )
    return o


class _Owned:
    """Sub-modules reach the owning Tacotron2 (its CUDA context) through a weak reference that is not part of their state:
    copies / pickles drop it and Tacotron2.__setstate__ re-links."""

    def __getstate__(self):
        st = dict(self.__dict__)
        st.pop("_t2_owner", None)
        return st


class Attention(_Owned, nn.Module):
    """model/attention.py:6-69."""

    def __init__(self, attention_rnn_dim, embedding_dim, attention_dim, attention_location_n_filters, attention_location_kernel_size):
        super().__init__()
        assert attention_location_n_filters == 32 and attention_location_kernel_size == 31, "kernels are specialised for 32 filters / k=31 (model/decoder.py:35-36)"
        self.query_layer = nn.Linear(attention_rnn_dim, attention_dim, bias=False)
        self.v = nn.Linear(attention_dim, 1, bias=False)
        self.location_conv = nn.Conv1d(2, attention_location_n_filters, kernel_size=attention_location_kernel_size,
                                       padding=int((attention_location_kernel_size - 1) / 2), bias=False, stride=1, dilation=1)
        self.location_dense = nn.Linear(attention_location_n_filters, attention_dim, bias=False)

    def forward(self, attention_hidden_state, memory, processed_memory, attention_weights_cat, mask):
        """model/attention.py:34-69 as a stand-alone call (C ABI ``t2_attention_step``).  Inside ``Decoder.forward`` and
        ``Tacotron2.forward`` the same arithmetic runs fused into the persistent decoder kernel."""
        return _owner_of(self)._attention_step(attention_hidden_state, memory, processed_memory, attention_weights_cat, mask)


class Decoder(_Owned, nn.Module):
    """model/decoder.py:10-119."""

    def __init__(self, num_mels, embedding_dim, prenet_dim, att_rnn_dim, att_dim, rnn_hidden_dim, dropout,
                 extra_att_in_dim: int = 0, extra_decoder_in_dim: int = 0):
        super().__init__()
        assert extra_att_in_dim == 0, "extra_att_in is never passed by the reference (model/tacotron2.py:297-308)"
        self.att_rnn = nn.LSTMCell(prenet_dim + embedding_dim + extra_att_in_dim, att_rnn_dim)
        self.att_rnn_dropout = nn.Dropout(0.1)
        self.attention = Attention(att_rnn_dim, embedding_dim, att_dim, 32, 31)
        self.lstm = nn.LSTMCell(att_rnn_dim + embedding_dim + extra_decoder_in_dim, rnn_hidden_dim)
        self.lstm_dropout = nn.Dropout(0.1)
        self.mel_out = nn.Linear(rnn_hidden_dim + embedding_dim + extra_decoder_in_dim, num_mels)
        self.gate = nn.Linear(rnn_hidden_dim + embedding_dim, 1)

    def forward(self, prev_mel_prenet, att_rnn_hidden, att_context, att_weights, att_weights_cum, rnn_hidden, encoded,
                att_encoded, encoded_mask, speech_features: Optional[Tensor] = None, extra_att_in: Optional[Tensor] = None,
                extra_decoder_in: Optional[Tensor] = None):
        assert speech_features is None and extra_att_in is None, "never passed by the reference"
        return _owner_of(self)._decoder_step(prev_mel_prenet, att_rnn_hidden, att_context, att_weights, att_weights_cum,
                                             rnn_hidden, encoded, att_encoded, encoded_mask, extra_decoder_in)


class Encoder(_Owned, nn.Module):
    """model/encoder.py:8-67."""

    def __init__(self, num_chars, embedding_dim, encoder_kernel_size, dropout):
        super().__init__()
        self.embedding = nn.Embedding(num_chars + 1, embedding_dim, padding_idx=0)
        self.embedding.weight.data.normal_(mean=0, std=0.5)
        convolutions = []
        for _ in range(3):
            convolutions.extend([
                nn.Conv1d(embedding_dim, embedding_dim, kernel_size=encoder_kernel_size, stride=1,
                          padding=int((encoder_kernel_size - 1) / 2), dilation=1),
                nn.BatchNorm1d(embedding_dim), nn.ReLU(), nn.Dropout(dropout),
            ])
        self.convolutions = nn.Sequential(*convolutions)
        self.lstm = nn.LSTM(embedding_dim, embedding_dim // 2, batch_first=True, bidirectional=True)

    def forward(self, char_idx, char_idx_len, dropout_keep=None):
        o = _owner_of(self)
        if self.training:
            return o._encode_train(char_idx, char_idx_len, dropout_keep)
        return o._encode(char_idx, char_idx_len, None)


class Postnet(_Owned, nn.Module):
    """model/postnet.py:4-52."""

    def __init__(self, num_layers, num_mels, postnet_dim, dropout):
        super().__init__()
        assert num_layers == 5, "the reference always builds 5 layers (model/tacotron2.py:122-124)"
        chans = [num_mels] + [postnet_dim] * (num_layers - 1) + [num_mels]
        layers = []
        for i in range(num_layers):
            layers.extend([nn.Conv1d(chans[i], chans[i + 1], 5, bias=False, padding="same"), nn.BatchNorm1d(chans[i + 1])])
            if i < num_layers - 1:
                layers.append(nn.Tanh())
            layers.append(nn.Dropout(dropout))
        self.postnet = nn.Sequential(*layers)

    def forward(self, X, dropout_keep=None):
        """X (B,M,T) -> (B,M,T), as the reference; internally channels-last.  ``dropout_keep``: optional (5,B,T,postnet_dim)
        uint8 keep-masks for the train-mode dropouts (parity tests)."""
        o = _owner_of(self)
        x = X.transpose(1, 2).contiguous()
        if self.training:
            from .autograd import PostnetFn
            keep = None if dropout_keep is None else dropout_keep.to(device=x.device, dtype=torch.uint8).contiguous()
            return PostnetFn.apply(o, keep, o._draw_seed(), x, *o._group_params("postnet").values()).transpose(1, 2)
        return o._postnet(x, x.shape[1], None).transpose(1, 2)


class Tacotron2(nn.Module):
    """model/tacotron2.py:14-347, same constructor and forward signature."""

    def __init__(self, num_chars: int, encoded_dim: int, encoder_kernel_size: int, num_mels: int, prenet_dim: int,
                 att_rnn_dim: int, att_dim: int, rnn_hidden_dim: int, postnet_dim: int, dropout: float,
                 speaker_tokens: bool = False, speaker_tokens_dim: Optional[int] = 128, num_speakers: int = 1,
                 controls: bool = False, controls_dim: int = 0, description_embeddings: bool = False,
                 description_embeddings_dim: int = 0):
        super().__init__()
        import weakref

        self.embedding_dim = encoded_dim
        self.num_mels = num_mels
        self.att_rnn_dim = att_rnn_dim
        self.rnn_hidden_dim = rnn_hidden_dim
        self.char_embedding_dim = encoded_dim
        self.prenet_dim = prenet_dim
        self.att_dim = att_dim
        self.postnet_dim = postnet_dim
        self.dropout = dropout
        self.controls = controls
        self.controls_dim = controls_dim if controls else 0
        self.speaker_tokens = speaker_tokens
        assert not speaker_tokens or num_speakers is not None, "If speaker tokens are enabled, you must give a num_speakers!"
        if speaker_tokens:  # model/tacotron2.py:61-65
            self.speaker_embedding = nn.Embedding(num_embeddings=num_speakers, embedding_dim=encoded_dim)
            self.speaker_embedding.weight.data.normal_(mean=0, std=0.5)
        self.description_embeddings = description_embeddings
        self.encoder = Encoder(num_chars=num_chars, embedding_dim=encoded_dim, encoder_kernel_size=encoder_kernel_size, dropout=dropout)
        self.prenet = nn.Sequential(  # model/tacotron2.py:85-92
            nn.Linear(num_mels, prenet_dim, bias=False), nn.ReLU(), AlwaysDropout(dropout),
            nn.Linear(prenet_dim, prenet_dim, bias=False), nn.ReLU(), AlwaysDropout(dropout),
        )
        encoded_full_dim = encoded_dim
        if description_embeddings:  # :98-104
            self.description_embeddings_linear = nn.Sequential(nn.Linear(description_embeddings_dim, 128), nn.Tanh())
            encoded_full_dim += 128
        self.encoded_full_dim = encoded_full_dim
        self.att_encoder = nn.Linear(encoded_full_dim, att_dim, bias=False)
        self.decoder = Decoder(num_mels=num_mels, embedding_dim=encoded_full_dim, prenet_dim=prenet_dim, att_rnn_dim=att_rnn_dim,
                               att_dim=att_dim, rnn_hidden_dim=rnn_hidden_dim, dropout=dropout,
                               extra_decoder_in_dim=self.controls_dim)
        self.postnet = Postnet(num_layers=5, num_mels=num_mels, postnet_dim=postnet_dim, dropout=dropout)

        self._cfg = _lib.T2ModelConfig(
            num_chars, encoded_dim, encoder_kernel_size, num_mels, prenet_dim, att_rnn_dim, att_dim, rnn_hidden_dim, postnet_dim,
            float(dropout), num_speakers if speaker_tokens else 0, self.controls_dim,
            description_embeddings_dim if description_embeddings else 0,
            1 if os.environ.get("T2B200_PRECISION", "fp32").lower() == "bf16" else 0)
        self._ctx = None
        self._sig = None
        self._ws = None
        ref = weakref.ref(self)
        for m in (self.encoder, self.decoder, self.decoder.attention, self.postnet):
            object.__setattr__(m, "_t2_owner", ref)

    # ------------------------------------------------------------------ context plumbing
    def __del__(self):
        try:
            if self._ctx is not None:
                _lib.load().t2_ctx_destroy(self._ctx)
        except Exception:
            pass

    # The CUDA context, its weight signature and the workspace are per-process device state: a copy / pickle of the module carries
    # the parameters only and rebuilds them lazily on first use (copy.deepcopy(model) / torch.save(model) used to raise).
    _TRANSIENT = ("_ctx", "_sig", "_ws", "_tape_pool", "_grad_bucket_hook", "_prof")

    def __getstate__(self):
        st = dict(self.__dict__)
        for k in self._TRANSIENT:
            st.pop(k, None)
        return st

    def __setstate__(self, st):
        import weakref
        self.__dict__.update(st)
        self._ctx = self._sig = self._ws = None
        ref = weakref.ref(self)
        for m in (self.encoder, self.decoder, self.decoder.attention, self.postnet):
            object.__setattr__(m, "_t2_owner", ref)

    def check_status(self):
        """Raises if a persistent kernel of this model aborted since the last check (C ABI ``t2_ctx_status``; synchronises)."""
        if self._ctx is None:
            return
        st = C.c_int(0)
        _lib.check(_lib.load().t2_ctx_status(self._ctx, C.byref(st)), "t2_ctx_status")
        if st.value != 0:
            raise RuntimeError("tacotron2_b200: a persistent decoder kernel aborted (grid barrier / pipeline timeout); "
                               "the outputs of that call were NaN-filled")

    def _tensors(self):
        out = dict(self.named_parameters())
        out.update({k: v for k, v in self.named_buffers() if v.is_floating_point()})
        return out

    def _ready(self, device):
        """Create the context on first use and re-pack weights whenever a parameter changed."""
        if device.type != "cuda":
            raise RuntimeError("tacotron2_b200 runs on CUDA (sm_100a) tensors only; there is no CPU fallback")
        lib = _lib.load()
        tensors = self._tensors()
        for k, v in tensors.items():
            if v.device != device or v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be a contiguous fp32 tensor on {device} (is {v.dtype} on {v.device})")
        sig = tuple((k, v.data_ptr(), v._version) for k, v in tensors.items())
        with torch.cuda.device(device):
            if self._ctx is None:
                h = C.c_void_p()
                _lib.check(lib.t2_ctx_create(C.byref(self._cfg), C.byref(h)), "t2_ctx_create")
                self._ctx = h
            if sig != self._sig or (getattr(self, "_bn_dirty", False) and not self.training):
                self._bn_dirty = False
                for k, v in tensors.items():
                    _lib.check(lib.t2_ctx_set_param(self._ctx, k.encode(), _lib.ptr(v), v.numel()), "t2_ctx_set_param")
                _lib.check(lib.t2_ctx_commit(self._ctx, self._stream()), "t2_ctx_commit")
                self._sig = sig
        return lib

    @staticmethod
    def _stream():
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def _workspace(self, lib, device, B, L, T):
        need = lib.t2_workspace_bytes(self._ctx, B, L, T)
        if need < 0:
            raise RuntimeError("t2_workspace_bytes failed")
        if self._ws is None or self._ws.numel() < need or self._ws.device != device:
            self._ws = torch.empty(need, dtype=torch.uint8, device=device)
        return self._ws

    _GROUPS = {
        "postnet": ("postnet.",),
        "decoder": ("decoder.", "prenet."),
        "memory": ("att_encoder.", "speaker_embedding.", "description_embeddings_linear."),
        "encoder": ("encoder.",),
    }

    def _group_params(self, group):
        """Parameters of one gradient bucket, in registration order (key -> Parameter)."""
        pre = self._GROUPS[group]
        return {k: p for k, p in self.named_parameters() if k.startswith(pre)}

    def _train_workspace(self, lib, device, B, L, T):
        need = lib.t2_train_workspace_bytes(self._ctx, B, L, T)
        if need < 0:
            raise RuntimeError("t2_train_workspace_bytes failed")
        if self._ws is None or self._ws.numel() < need or self._ws.device != device:
            self._ws = torch.empty(need, dtype=torch.uint8, device=device)
        return self._ws

    @staticmethod
    def _draw_seed():
        return int(torch.randint(0, 2**62, (1,)).item())  # CPU generator: follows torch.manual_seed like run/say.py:40

    def _decode_tf_train(self, memory, pmem, mask, ctrl, target, masks):
        """Teacher-forced frame loop under train() with autograd (<= 64 utterances per launch; larger batches are chunked --
        utterances are independent inside the loop, and autograd sums the per-chunk parameter gradients)."""
        from .autograd import DecodeTfFn
        dev = memory.device
        B = memory.shape[0]
        pk, ak, rk = (self._u8(masks.get(k), dev) for k in ("prenet", "att_h", "rnn_h"))
        params = list(self._group_params("decoder").values())
        seed = self._draw_seed()
        outs = []
        for b0 in range(0, B, _MAX_DECODE_BATCH):
            b1 = min(B, b0 + _MAX_DECODE_BATCH)
            whole = b0 == 0 and b1 == B
            sl = (lambda t: t) if whole else (lambda t: None if t is None else t[b0:b1].contiguous())
            sk = (lambda t: t) if whole else (lambda t: None if t is None else t[:, :, b0:b1].contiguous())
            sh = (lambda t: t) if whole else (lambda t: None if t is None else t[:, b0:b1].contiguous())
            outs.append(DecodeTfFn.apply(self, sl(mask), sl(ctrl), sl(target), sk(pk), sh(ak), sh(rk), seed + b0, sl(memory), sl(pmem), *params))
        if len(outs) == 1:
            return outs[0]
        return tuple(torch.cat([o[i] for o in outs], 0) for i in range(3))

    def loss(self, mels, mels_post, gates, mel_target, gate_target):
        """training_step loss (model/tts_model.py:197-201) as one kernel: BCEWithLogits + MSE + MSE."""
        from .autograd import LossFn
        return LossFn.apply(self, mels, mels_post, gates, mel_target, gate_target)

    @staticmethod
    def _u8(t, device):
        return None if t is None else t.to(device=device, dtype=torch.uint8).contiguous()

    # ------------------------------------------------------------------ stages
    def _encode(self, chars_idx, chars_idx_len, keep):
        lib = self._ready(chars_idx.device)
        B, L = chars_idx.shape
        dev = chars_idx.device
        ws = self._workspace(lib, dev, B, L, 1)
        chars = chars_idx.to(torch.int64).contiguous()
        lens = chars_idx_len.to(device=dev, dtype=torch.int64).contiguous()
        out = torch.empty(B, L, self.embedding_dim, device=dev)
        train = 1 if self.training else 0
        _lib.check(lib.t2_encoder_fwd(self._ctx, _lib.ptr(chars), _lib.ptr(lens), B, L, train, _lib.ptr(keep), _lib.ptr(out),
                                      _lib.ptr(ws), ws.numel(), self._stream()), "t2_encoder_fwd")
        return out

    def _encode_train(self, chars_idx, chars_idx_len, keep):
        from .autograd import EncoderFn
        dev = chars_idx.device
        chars = chars_idx.to(torch.int64).contiguous()
        lens = chars_idx_len.to(device=dev, dtype=torch.int64).contiguous()
        return EncoderFn.apply(self, chars, lens, self._u8(keep, dev), self._draw_seed(), *self._group_params("encoder").values())

    def _forward_train(self, chars_idx, chars_idx_len, mel_spectrogram, mel_spectrogram_len, speaker_id, controls, max_len,
                       description_embeddings, masks):
        """Teacher-forced forward under train() (what training_step runs, model/tts_model.py:186-195) with autograd."""
        from .autograd import FinalizeFn, MemoryFn, PostnetFn
        dev = chars_idx.device
        B, L = chars_idx.shape
        encoded = self._encode_train(chars_idx, chars_idx_len, masks.get("encoder"))
        spk = speaker_id.to(device=dev, dtype=torch.int64).contiguous() if self.speaker_tokens else None
        desc = description_embeddings.to(device=dev, dtype=torch.float32).contiguous() if self.description_embeddings else None
        memory, pmem = MemoryFn.apply(self, spk, desc, encoded, *self._group_params("memory").values())
        lens = chars_idx_len.to(device=dev, dtype=torch.int64)
        mask = (torch.arange(L, device=dev)[None, :] >= lens[:, None]).to(torch.uint8).contiguous()
        ctrl = controls.to(device=dev, dtype=torch.float32).contiguous() if self.controls else None
        target = mel_spectrogram.to(device=dev, dtype=torch.float32)[:, :max_len].contiguous()
        mels_raw, gates_raw, aligns = self._decode_tf_train(memory, pmem, mask, ctrl, target, masks)
        post_keep = self._u8(masks.get("postnet"), dev)
        post_out = PostnetFn.apply(self, post_keep, self._draw_seed(), mels_raw, *self._group_params("postnet").values())
        lengths = mel_spectrogram_len.to(device=dev, dtype=torch.int64).contiguous()
        mels, post, gates = FinalizeFn.apply(self, lengths, mels_raw, post_out, gates_raw)
        return mels, post, gates.unsqueeze(-1), aligns

    def _memory(self, encoded, speaker_id, description):
        lib = self._ready(encoded.device)
        B, L, _ = encoded.shape
        dev = encoded.device
        ws = self._workspace(lib, dev, B, L, 1)
        memory = torch.empty(B, L, self.encoded_full_dim, device=dev)
        pmem = torch.empty(B, L, self.att_dim, device=dev)
        spk = None if speaker_id is None else speaker_id.to(device=dev, dtype=torch.int64).contiguous()
        desc = None if description is None else description.to(device=dev, dtype=torch.float32).contiguous()
        enc_c = encoded.contiguous()   # bound to a local until after the launch
        _lib.check(lib.t2_memory_fwd(self._ctx, _lib.ptr(enc_c), _lib.ptr(spk), _lib.ptr(desc), B, L, _lib.ptr(memory),
                                     _lib.ptr(pmem), _lib.ptr(ws), ws.numel(), self._stream()), "t2_memory_fwd")
        return memory, pmem

    def _decode_ar_chunked(self, lib, memory, pmem, mask, ctrl, B, L, max_len, pk, seed, mels, gates, aligns, lengths, ws, st):
        """Autoregressive decode of more than 64 utterances.  The reference's loop ends when EVERY utterance of the batch is done
        (model/tacotron2.py:321-322) and finished utterances keep decoding until then (:319-325), so the chunks (<= 64 utterances
        per persistent-kernel launch) must all produce the frame count of the slowest one: every chunk is decoded once with its own
        stop condition; chunks that stopped early are decoded again with the batch's frame count as the minimum (the kernel is
        deterministic and the dropout draws depend only on (seed, frame, utterance), so the prefix is reproduced bit for bit)."""
        dev = memory.device
        chunks = [(b0, min(B, b0 + _MAX_DECODE_BATCH)) for b0 in range(0, B, _MAX_DECODE_BATCH)]
        nf = torch.zeros(len(chunks), dtype=torch.int32, device=dev)

        def run(i, min_frames):
            b0, b1 = chunks[i]
            pkc = None if pk is None else pk[:, :, b0:b1].contiguous()
            cc = None if ctrl is None else ctrl[b0:b1]
            _lib.check(lib.t2_decode_ar_chunk(self._ctx, _lib.ptr(memory[b0:b1]), _lib.ptr(pmem[b0:b1]), _lib.ptr(mask[b0:b1]), _lib.ptr(cc),
                                              b1 - b0, L, max_len, min_frames, _lib.ptr(pkc), seed + b0, _lib.ptr(mels[b0:b1]),
                                              _lib.ptr(gates[b0:b1]), _lib.ptr(aligns[b0:b1]), _lib.ptr(lengths[b0:b1]), _lib.ptr(nf[i:i + 1]),
                                              _lib.ptr(ws), ws.numel(), st), "t2_decode_ar_chunk")
            return pkc   # keep the sliced masks alive until the stream has consumed them

        keep = [run(i, 0) for i in range(len(chunks))]
        n_c = nf.tolist()   # one host sync for the whole batch
        if min(n_c) < 0:
            raise RuntimeError("decoder kernel aborted (grid barrier timeout)")
        n = max(n_c)
        keep += [run(i, n) for i in range(len(chunks)) if n_c[i] < n]
        n_c = nf.tolist()
        if min(n_c) < 0:
            raise RuntimeError("decoder kernel aborted (grid barrier timeout)")
        assert all(x == n for x in n_c), (n_c, n)
        del keep
        return n

    def _postnet(self, mels_cl, t_stride, keep):
        """mels_cl: (B,T,M) view whose batch stride is t_stride*M."""
        lib = self._ready(mels_cl.device)
        B, T, M = mels_cl.shape
        dev = mels_cl.device
        ws = self._workspace(lib, dev, B, 1, T)
        out = torch.empty(B, T, M, device=dev)
        _lib.check(lib.t2_postnet_fwd(self._ctx, _lib.ptr(mels_cl), t_stride, B, T, 1 if self.training else 0, _lib.ptr(keep),
                                      _lib.ptr(out), _lib.ptr(ws), ws.numel(), self._stream()), "t2_postnet_fwd")
        return out

    def _decoder_step(self, prev, att_hidden, att_context, w, w_cum, rnn_hidden, encoded, att_encoded, mask, controls):
        lib = self._ready(prev.device)
        dev = prev.device
        B, L = w.shape
        ws = self._workspace(lib, dev, B, L, 1)
        f = lambda t: t.to(torch.float32).contiguous()  # noqa: E731
        e = lambda *s: torch.empty(*s, device=dev)  # noqa: E731
        H, R, EF, M = self.att_rnn_dim, self.rnn_hidden_dim, self.encoded_full_dim, self.num_mels
        mel, gate, ah, ac, ctx = e(B, M), e(B, 1), e(B, H), e(B, H), e(B, EF)
        w2, wc2, rh, rc = e(B, L), e(B, L), e(B, R), e(B, R)
        # converted inputs are bound to locals that live until after the call: a temporary freed before the launch could hand its
        # block to the next conversion
        ins = [f(prev), f(att_hidden[0]), f(att_hidden[1]), f(att_context), f(w), f(w_cum), f(rnn_hidden[0]), f(rnn_hidden[1]),
               f(encoded), f(att_encoded), self._u8(mask, dev), None if controls is None else f(controls)]
        _lib.check(lib.t2_decoder_step(
            self._ctx, *[_lib.ptr(t) for t in ins], B, L,
            _lib.ptr(mel), _lib.ptr(gate), _lib.ptr(ah), _lib.ptr(ac), _lib.ptr(ctx), _lib.ptr(w2), _lib.ptr(wc2), _lib.ptr(rh),
            _lib.ptr(rc), _lib.ptr(ws), ws.numel(), self._stream()), "t2_decoder_step")
        del ins
        w_cum.copy_(wc2)  # the reference updates the caller's tensor in place (model/decoder.py:90)
        return mel, gate, (ah, ac), ctx, w2, w_cum, (rh, rc)

    def _attention_step(self, attention_hidden_state, memory, processed_memory, attention_weights_cat, mask):
        dev = attention_hidden_state.device
        lib = self._ready(dev)
        B, L = processed_memory.shape[0], processed_memory.shape[1]
        f = lambda t: t.to(device=dev, dtype=torch.float32).contiguous()  # noqa: E731
        ins = [f(attention_hidden_state), f(memory), f(processed_memory), f(attention_weights_cat), self._u8(mask, dev)]
        assert tuple(ins[3].shape) == (B, 2, L) and tuple(ins[4].shape) == (B, L), "attention_weights_cat (B,2,L), mask (B,L)"
        context = torch.empty(B, self.encoded_full_dim, device=dev)
        weights = torch.empty(B, L, device=dev)
        _lib.check(lib.t2_attention_step(self._ctx, *[_lib.ptr(t) for t in ins], B, L, _lib.ptr(context), _lib.ptr(weights),
                                         self._stream()), "t2_attention_step")
        del ins
        return context, weights

    def init_hidden(self, encoded_len: int, batch_size: int, device: torch.device):
        """model/tacotron2.py:126-153."""
        z = lambda *s: torch.zeros(*s, device=device)  # noqa: E731
        return ((z(batch_size, self.att_rnn_dim), z(batch_size, self.att_rnn_dim)), z(batch_size, self.encoded_full_dim),
                z(batch_size, encoded_len), z(batch_size, encoded_len),
                (z(batch_size, self.rnn_hidden_dim), z(batch_size, self.rnn_hidden_dim)))

    # ------------------------------------------------------------------ forward
    def forward(self, chars_idx: Tensor, chars_idx_len: Tensor, teacher_forcing: bool, mel_spectrogram: Optional[Tensor] = None,
                mel_spectrogram_len: Optional[Tensor] = None, speaker_id: Optional[Tensor] = None,
                controls: Optional[Tensor] = None, max_len_override: Optional[int] = None,
                description_embeddings: Optional[Tensor] = None, dropout_masks: Optional[dict] = None):
        # argument checks: same conditions and messages as model/tacotron2.py:167-189,247-250
        if teacher_forcing:
            assert mel_spectrogram is not None, "Ground-truth Mel spectrogram is required for teacher forcing"
            assert mel_spectrogram_len is not None, "Ground-truth Mel spectrogram lengths are required for teacher forcing"
        assert not self.speaker_tokens or speaker_id is not None, "speaker_id tensor required when speaker tokens are active!"
        assert not self.description_embeddings or description_embeddings is not None, \
            "description tensor required when description tokens are active!"
        assert not self.controls or controls is not None, "Controls are enabled, but no control vector was passed to the model!"
        assert self.controls or controls is None, "Controls are disabled, but a control vector was passed to the model!"
        if max_len_override is None and mel_spectrogram is None:
            raise Exception("If Mel spectrogram is not given, max_len_override is required!")

        dev = chars_idx.device
        lib = self._ready(dev)
        B, L = chars_idx.shape
        masks = dropout_masks or {}
        train = bool(self.training)
        max_len = max_len_override if max_len_override is not None else mel_spectrogram.shape[1]

        if train:
            if not teacher_forcing:
                raise RuntimeError("train() mode supports the teacher-forced forward only (training_step, model/tts_model.py:186-195); "
                                   "call eval() for autoregressive synthesis")
            return self._forward_train(chars_idx, chars_idx_len, mel_spectrogram, mel_spectrogram_len, speaker_id, controls, max_len,
                                       description_embeddings, masks)
        encoded = self._encode(chars_idx, chars_idx_len, None)
        memory, pmem = self._memory(encoded, speaker_id if self.speaker_tokens else None,
                                    description_embeddings if self.description_embeddings else None)
        lens = chars_idx_len.to(device=dev, dtype=torch.int64)
        mask = (torch.arange(L, device=dev)[None, :] >= lens[:, None]).to(torch.uint8).contiguous()  # :223-226
        ctrl = controls.to(device=dev, dtype=torch.float32).contiguous() if self.controls else None
        pk = self._u8(masks.get("prenet"), dev)
        if pk is not None:
            assert tuple(pk.shape) == (pk.shape[0], 2, B, self.prenet_dim) and pk.shape[0] >= max_len + (1 if teacher_forcing else 0), \
                f"prenet keep-mask must be (>= T+1, 2, B, P), got {tuple(pk.shape)}"
        seed = int(torch.randint(0, 2**62, (1,)).item())  # CPU generator: follows torch.manual_seed like run/say.py:40

        M = self.num_mels
        mels = torch.empty(B, max_len, M, device=dev)
        gates = torch.empty(B, max_len, device=dev)
        aligns = torch.empty(B, max_len, L, device=dev)
        ws = self._workspace(lib, dev, min(B, _MAX_DECODE_BATCH), L, max_len)
        st = self._stream()
        if teacher_forcing:
            T = max_len
            assert T <= mel_spectrogram.shape[1]
            target = mel_spectrogram.to(device=dev, dtype=torch.float32)
            ak = self._u8(masks.get("att_h"), dev) if train else None
            rk = self._u8(masks.get("rnn_h"), dev) if train else None
            for b0 in range(0, B, _MAX_DECODE_BATCH):  # utterances are independent: chunks of <= 64 are exact
                b1 = min(B, b0 + _MAX_DECODE_BATCH)
                whole = b0 == 0 and b1 == B
                sl = (lambda t: t) if whole else (lambda t: None if t is None else t[b0:b1].contiguous())
                sk = (lambda t: t) if whole else (lambda t: None if t is None else t[:, :, b0:b1].contiguous())
                sh = (lambda t: t) if whole else (lambda t: None if t is None else t[:, b0:b1].contiguous())
                m_c, g_c, a_c = (mels, gates, aligns) if whole else (torch.empty(b1 - b0, T, M, device=dev),
                                                                     torch.empty(b1 - b0, T, device=dev),
                                                                     torch.empty(b1 - b0, T, L, device=dev))
                tgt = sl(target[:, :T].contiguous())
                _lib.check(lib.t2_decode_tf(self._ctx, _lib.ptr(sl(memory)), _lib.ptr(sl(pmem)), _lib.ptr(sl(mask)), _lib.ptr(sl(ctrl)),
                                            _lib.ptr(tgt), b1 - b0, L, T, _lib.ptr(sk(pk)), 1 if train else 0, _lib.ptr(sh(ak)),
                                            _lib.ptr(sh(rk)), seed + b0, _lib.ptr(m_c), _lib.ptr(g_c), _lib.ptr(a_c), _lib.ptr(ws),
                                            ws.numel(), st), "t2_decode_tf")
                if not whole:
                    mels[b0:b1], gates[b0:b1], aligns[b0:b1] = m_c, g_c, a_c
            n = T
            lengths = mel_spectrogram_len.to(device=dev, dtype=torch.int64).contiguous()  # :271
        else:
            lengths = torch.empty(B, dtype=torch.int64, device=dev)
            if B > _MAX_DECODE_BATCH:
                n = self._decode_ar_chunked(lib, memory, pmem, mask, ctrl, B, L, max_len, pk, seed, mels, gates, aligns, lengths, ws, st)
                post_keep = self._u8(masks.get("postnet"), dev) if train else None
                post = self._postnet(mels[:, :n], max_len, post_keep)
                _lib.check(lib.t2_finalize(self._ctx, _lib.ptr(mels), _lib.ptr(post), _lib.ptr(gates), _lib.ptr(lengths), B, n, max_len, st),
                           "t2_finalize")
                return mels[:, :n], post, gates[:, :n].unsqueeze(-1), aligns[:, :n]
            nf = torch.zeros(1, dtype=torch.int32, device=dev)
            prof = getattr(self, "_prof", None)  # bench.py: CUDA events around the persistent kernel, same stream
            if prof is not None:
                ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                ev[0].record()
            _lib.check(lib.t2_decode_ar(self._ctx, _lib.ptr(memory), _lib.ptr(pmem), _lib.ptr(mask), _lib.ptr(ctrl), B, L, max_len,
                                        _lib.ptr(pk), seed, _lib.ptr(mels), _lib.ptr(gates), _lib.ptr(aligns), _lib.ptr(lengths),
                                        _lib.ptr(nf), _lib.ptr(ws), ws.numel(), st), "t2_decode_ar")
            if prof is not None:
                ev[1].record()
                prof.append(ev)
            n = int(nf.item())  # one host sync per utterance batch (the reference syncs every frame, :321)
            if n < 0:
                self.check_status()
                raise RuntimeError("decoder kernel aborted (grid barrier timeout)")
        post_keep = self._u8(masks.get("postnet"), dev) if train else None
        post = self._postnet(mels[:, :n], max_len, post_keep)
        _lib.check(lib.t2_finalize(self._ctx, _lib.ptr(mels), _lib.ptr(post), _lib.ptr(gates), _lib.ptr(lengths), B, n, max_len, st),
                   "t2_finalize")
        return mels[:, :n], post, gates[:, :n].unsqueeze(-1), aligns[:, :n]
