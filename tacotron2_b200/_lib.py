"""ctypes binding of libt2b200.so (C ABI declared in include/t2b200.h).

The library is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  There is no
fallback of any kind: if the shared object is missing or a call fails, this raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("T2B200_LIB") or os.path.join(_HERE, "libt2b200.so")   # (T2B200_LIB: A/B builds of the same library, tools/)


class T2ModelConfig(C.Structure):
    _fields_ = [
        ("num_chars", C.c_int), ("encoded_dim", C.c_int), ("encoder_kernel_size", C.c_int), ("num_mels", C.c_int),
        ("prenet_dim", C.c_int), ("att_rnn_dim", C.c_int), ("att_dim", C.c_int), ("rnn_hidden_dim", C.c_int),
        ("postnet_dim", C.c_int), ("dropout", C.c_float), ("num_speakers", C.c_int), ("controls_dim", C.c_int),
        ("description_dim", C.c_int), ("precision", C.c_int),
    ]


class T2HifiganConfig(C.Structure):
    _fields_ = [
        ("num_mels", C.c_int), ("upsample_initial_channel", C.c_int), ("num_upsamples", C.c_int), ("upsample_rates", C.c_int * 8),
        ("upsample_kernel_sizes", C.c_int * 8), ("resblock", C.c_int), ("num_kernels", C.c_int), ("resblock_kernel_sizes", C.c_int * 8),
        ("resblock_dilation_sizes", (C.c_int * 3) * 8), ("precision", C.c_int),
    ]


class T2BertConfig(C.Structure):
    _fields_ = [("vocab_size", C.c_int), ("hidden_size", C.c_int), ("num_layers", C.c_int), ("num_heads", C.c_int), ("intermediate_size", C.c_int),
                ("max_position", C.c_int), ("type_vocab_size", C.c_int), ("layer_norm_eps", C.c_float), ("precision", C.c_int)]


_P, _I, _LL, _ULL = C.c_void_p, C.c_int, C.c_longlong, C.c_ulonglong

# name -> (restype, argtypes); kept in sync with include/t2b200.h (tests/test_abi.py checks the symbol list)
SIGNATURES = {
    "t2_last_error": (C.c_char_p, []),
    "t2_abi_version": (_I, []),
    "t2_ctx_create": (_I, [C.POINTER(T2ModelConfig), C.POINTER(_P)]),
    "t2_ctx_destroy": (_I, [_P]),
    "t2_ctx_set_param": (_I, [_P, C.c_char_p, _P, _LL]),
    "t2_ctx_commit": (_I, [_P, _P]),
    "t2_workspace_bytes": (_LL, [_P, _I, _I, _I]),
    "t2_encoder_fwd": (_I, [_P, _P, _P, _I, _I, _I, _P, _P, _P, _LL, _P]),
    "t2_memory_fwd": (_I, [_P, _P, _P, _P, _I, _I, _P, _P, _P, _LL, _P]),
    "t2_decode_ar": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P, _ULL, _P, _P, _P, _P, _P, _P, _LL, _P]),
    "t2_decode_ar_chunk": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P, _ULL, _P, _P, _P, _P, _P, _P, _LL, _P]),
    "t2_decode_tf": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P, _I, _P, _P, _ULL, _P, _P, _P, _P, _LL, _P]),
    "t2_decoder_step": (_I, [_P] + [_P] * 12 + [_I, _I] + [_P] * 9 + [_P, _LL, _P]),
    "t2_attention_step": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P]),
    "t2_ctx_status": (_I, [_P, C.POINTER(_I)]),
    "t2_debug_set_timing": (_I, [_P, _I]),
    "t2_debug_kernel_ms": (_I, [_P, _I, C.POINTER(C.c_float)]),
    "t2_hifigan_create": (_I, [C.POINTER(T2HifiganConfig), C.POINTER(_P)]),
    "t2_hifigan_destroy": (_I, [_P]),
    "t2_hifigan_set_param": (_I, [_P, C.c_char_p, _P, _LL]),
    "t2_hifigan_commit": (_I, [_P, _P]),
    "t2_hifigan_workspace_bytes": (_LL, [_P, _I, _I]),
    "t2_hifigan_out_len": (_LL, [_P, _I]),
    "t2_hifigan_fwd": (_I, [_P, _P, _I, _I, _P, _P, _LL, _P]),
    "t2_bert_create": (_I, [C.POINTER(T2BertConfig), C.POINTER(_P)]),
    "t2_bert_destroy": (_I, [_P]),
    "t2_bert_set_param": (_I, [_P, C.c_char_p, _P, _LL]),
    "t2_bert_commit": (_I, [_P, _P]),
    "t2_bert_workspace_bytes": (_LL, [_P, _I, _I]),
    "t2_bert_pooler_fwd": (_I, [_P, _P, _P, _P, _I, _I, _P, _P, _P, _LL, _P]),
    "t2_prenet_fwd": (_I, [_P, _P, _P, _P, _ULL, _I, _P, _P, _LL, _P]),
    "t2_postnet_fwd": (_I, [_P, _P, _I, _I, _I, _I, _P, _P, _P, _LL, _P]),
    "t2_debug_phase_cycles": (_I, [_P, _P]),
    "t2_debug_tcgemm": (_I, [_P, _P, _P, _I, _I, _I, _P, _LL, _P]),
    "t2_finalize": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "t2_ctx_set_grad": (_I, [_P, C.c_char_p, _P, _LL]),
    "t2_train_workspace_bytes": (_LL, [_P, _I, _I, _I]),
    "t2_postnet_tape_bytes": (_LL, [_P, _I, _I]),
    "t2_postnet_train_fwd": (_I, [_P, _P, _I, _I, _I, _P, _ULL, _P, _P, _LL, _P, _LL, _P]),
    "t2_postnet_bwd": (_I, [_P, _I, _I, _P, _ULL, _P, _P, _P, _LL, _P, _P, _LL, _P]),
    "t2_encoder_tape_bytes": (_LL, [_P, _I, _I]),
    "t2_encoder_train_fwd": (_I, [_P, _P, _P, _I, _I, _P, _ULL, _P, _P, _LL, _P, _LL, _P]),
    "t2_encoder_bwd": (_I, [_P, _P, _P, _I, _I, _P, _ULL, _P, _P, _P, _LL, _P, _LL, _P]),
    "t2_memory_bwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _LL, _P]),
    "t2_decode_tape_bytes": (_LL, [_P, _I, _I, _I]),
    "t2_decode_tf_train_fwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _P, _ULL, _P, _P, _P, _P, _LL, _P, _LL, _P]),
    "t2_decode_tf_bwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _ULL, _P, _P, _P, _P, _LL, _P, _P, _P, _LL, _P]),
    "t2_loss_fwd_bwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P, _P, _LL, _P]),
    "t2_finalize_bwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _P, _P, _P, _P]),
    "t2_melspec": (_I, [_P, _P, _I, _LL, _P, _P, _P, _I, _I, _I, _P, _I, _P, _P]),
    "t2_trim_workspace_bytes": (_LL, [_I, _LL, _I]),
    "t2_trim": (_I, [_P, _P, _I, _LL, _LL, C.c_float, _I, _I, _P, _P, _P, _LL, _P]),
    "t2_trim_static": (_I, [_P, _P, _I, _LL, _LL, C.c_float, _I, _I, _P, _P, _P]),
    "t2_collate": (_I, [_I, _P, _P, _P, _P, _P, _I, _P, _P, _LL, _P]),
    "t2_clip_adam_chunks": (_LL, [_I, _P]),
    "t2_clip_adam_workspace_bytes": (_LL, [_I, _LL]),
    "t2_clip_adam_step": (_I, [_I, _P, _P, _P, _P, _P, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _LL, _I, _P, _P,
                               _LL, _P]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`. "
            "tacotron2_b200 has no CPU or PyTorch fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is missing
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def check(rc: int, what: str = "libt2b200") -> None:
    if rc != 0:
        msg = load().t2_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def ptr(t):
    """Device pointer of a torch tensor (or None)."""
    return None if t is None else C.c_void_p(t.data_ptr())
