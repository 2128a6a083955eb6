"""Description-embedding front end: ``BertModel(...).pooler_output`` on the GPU through libt2b200.

The reference turns a free-text voice description into a 768-d vector with Hugging Face's ``BertModel`` (``run/say.py:93-106``:
``AutoTokenizer`` + ``BertModel.from_pretrained("google-bert/bert-base-uncased")(...).pooler_output``) and trains from per-item
``.pt`` files holding such vectors (``datasets/tts_dataset.py:253-291``).  ``BertPooler`` keeps ``BertModel``'s parameter names
(a ``BertModel.state_dict()`` loads with ``load_state_dict``; ``from_hf`` copies from a live model) and the call signature
``(input_ids, token_type_ids=None, attention_mask=None)`` returning an object with ``pooler_output`` / ``last_hidden_state``.
The ``torch.nn`` sub-modules are parameter containers; the arithmetic runs in ``t2_bert_pooler_fwd`` (tcgen05 dense layers,
``csrc/bert.cu``).  Tokenisation stays with the caller's tokenizer (WordPiece needs the vocabulary file; nothing here ships it).
CUDA tensors only; no fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from types import SimpleNamespace

import torch
from torch import nn

from . import _lib

BERT_BASE = dict(vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
                 max_position_embeddings=512, type_vocab_size=2, layer_norm_eps=1e-12)


def _cfg(config, k):
    return config[k] if isinstance(config, dict) else getattr(config, k)


class _Self(nn.Module):
    def __init__(self, h):
        super().__init__()
        self.query, self.key, self.value = nn.Linear(h, h), nn.Linear(h, h), nn.Linear(h, h)


class _DenseLN(nn.Module):
    def __init__(self, i, o, eps):
        super().__init__()
        self.dense, self.LayerNorm = nn.Linear(i, o), nn.LayerNorm(o, eps=eps)


class _Attention(nn.Module):
    def __init__(self, h, eps):
        super().__init__()
        self.self, self.output = _Self(h), _DenseLN(h, h, eps)


class _Dense(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.dense = nn.Linear(i, o)


class _Layer(nn.Module):
    def __init__(self, h, inter, eps):
        super().__init__()
        self.attention, self.intermediate, self.output = _Attention(h, eps), _Dense(h, inter), _DenseLN(inter, h, eps)


class _Embeddings(nn.Module):
    def __init__(self, v, h, p, t, eps):
        super().__init__()
        self.word_embeddings, self.position_embeddings, self.token_type_embeddings = nn.Embedding(v, h), nn.Embedding(p, h), nn.Embedding(t, h)
        self.LayerNorm = nn.LayerNorm(h, eps=eps)


class _Encoder(nn.Module):
    def __init__(self, n, h, inter, eps):
        super().__init__()
        self.layer = nn.ModuleList([_Layer(h, inter, eps) for _ in range(n)])


class BertPooler(nn.Module):
    def __init__(self, config=None):
        super().__init__()
        config = BERT_BASE if config is None else config
        g = lambda k: _cfg(config, k)  # noqa: E731
        h, eps = int(g("hidden_size")), float(g("layer_norm_eps"))
        self.embeddings = _Embeddings(int(g("vocab_size")), h, int(g("max_position_embeddings")), int(g("type_vocab_size")), eps)
        self.encoder = _Encoder(int(g("num_hidden_layers")), h, int(g("intermediate_size")), eps)
        self.pooler = _Dense(h, h)
        self.hidden_size = h
        self._cfg = _lib.T2BertConfig(int(g("vocab_size")), h, int(g("num_hidden_layers")), int(g("num_attention_heads")), int(g("intermediate_size")),
                                      int(g("max_position_embeddings")), int(g("type_vocab_size")), eps,
                                      1 if os.environ.get("T2B200_PRECISION", "fp32").lower() == "bf16" else 0)
        self._ctx, self._sig, self._ws = None, None, None

    @classmethod
    def from_hf(cls, hf_model):
        """Copy the weights of a ``transformers.BertModel`` (``run/say.py:99``)."""
        m = cls(hf_model.config)
        m.load_state_dict({k: v for k, v in hf_model.state_dict().items() if not k.endswith("position_ids")})
        return m

    def __del__(self):
        try:
            if self._ctx is not None:
                _lib.load().t2_bert_destroy(self._ctx)
        except Exception:
            pass

    def _ready(self, device):
        if device.type != "cuda":
            raise RuntimeError("tacotron2_b200.bert runs on CUDA (sm_100a) tensors only; there is no CPU fallback")
        lib = _lib.load()
        tensors = dict(self.named_parameters())
        for k, v in tensors.items():
            if v.device != device or v.dtype != torch.float32 or not v.is_contiguous():
                raise RuntimeError(f"parameter {k} must be a contiguous fp32 tensor on {device}")
        sig = tuple((k, v.data_ptr(), v._version) for k, v in tensors.items())
        with torch.cuda.device(device):
            if self._ctx is None:
                hdl = C.c_void_p()
                _lib.check(lib.t2_bert_create(C.byref(self._cfg), C.byref(hdl)), "t2_bert_create")
                self._ctx = hdl
            if sig != self._sig:
                for k, v in tensors.items():
                    _lib.check(lib.t2_bert_set_param(self._ctx, k.encode(), _lib.ptr(v), v.numel()), "t2_bert_set_param")
                _lib.check(lib.t2_bert_commit(self._ctx, C.c_void_p(torch.cuda.current_stream().cuda_stream)), "t2_bert_commit")
                self._sig = sig
        return lib

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, output_hidden=True):
        dev = input_ids.device
        lib = self._ready(dev)
        B, S = input_ids.shape
        ids = input_ids.to(torch.int64).contiguous()
        tt = None if token_type_ids is None else token_type_ids.to(device=dev, dtype=torch.int64).contiguous()
        am = None if attention_mask is None else attention_mask.to(device=dev, dtype=torch.float32).contiguous()
        need = lib.t2_bert_workspace_bytes(self._ctx, B, S)
        if self._ws is None or self._ws.numel() < need or self._ws.device != dev:
            self._ws = None
            self._ws = torch.empty(need, dtype=torch.uint8, device=dev)
        pooled = torch.empty(B, self.hidden_size, device=dev)
        hidden = torch.empty(B, S, self.hidden_size, device=dev) if output_hidden else None
        _lib.check(lib.t2_bert_pooler_fwd(self._ctx, _lib.ptr(ids), _lib.ptr(tt), _lib.ptr(am), B, S, _lib.ptr(pooled), _lib.ptr(hidden),
                                          _lib.ptr(self._ws), self._ws.numel(), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                   "t2_bert_pooler_fwd")
        return SimpleNamespace(pooler_output=pooled, last_hidden_state=hidden)

    def describe(self, texts, tokenizer):
        """``run/say.py:93-106`` for a batch of description strings: ``tokenizer`` is the caller's ``AutoTokenizer`` (padding=True)."""
        tok = tokenizer(list(texts), return_tensors="pt", padding=True)
        dev = next(self.parameters()).device
        return self(tok["input_ids"].to(dev), tok.get("token_type_ids"), tok.get("attention_mask"), output_hidden=False).pooler_output


class DescriptionEmbeddingCache:
    """The per-item description files the reference's dataset reads (``datasets/tts_dataset.py:253-291``:
    ``torch.load(base_dir/<description_embeddings[i]>).unsqueeze(0)``): one ``.pt`` holding a (dim,) CPU tensor per item."""

    def __init__(self, base_dir: str):
        self.base_dir = base_dir

    def path(self, rel: str) -> str:
        return os.path.join(self.base_dir, rel)

    def save_batch(self, rels, embeddings) -> None:
        emb = embeddings.detach().to("cpu", torch.float32)
        for r, e in zip(rels, emb):
            os.makedirs(os.path.dirname(self.path(r)) or ".", exist_ok=True)
            torch.save(e.clone(), self.path(r))

    def load_batch(self, rels, dim: int, device=None):
        """(B, dim): missing / None entries are zeros, as the dataset does (:287-290)."""
        rows = [torch.zeros(dim) if r is None or not os.path.exists(self.path(r)) else torch.load(self.path(r), map_location="cpu") for r in rels]
        out = torch.stack(rows)
        return out if device is None else out.to(device)
