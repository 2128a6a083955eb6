"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports exactly the symbols
include/t2b200.h declares, and the host mirror keeps the reference's module API (no GPU needed)."""
import inspect
import os
import re

import pytest
import torch

import __graft_entry__ as G
from tacotron2_b200 import _lib, synth
from tacotron2_b200.model import AlwaysDropout, Attention, Decoder, Encoder, Postnet, Tacotron2

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    G.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "t2b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(t2_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.t2_abi_version() == 1


def test_no_cpu_fallback(lib):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import ctypes as C
    cfg = _lib.T2ModelConfig(39, 512, 5, 80, 256, 1024, 128, 1024, 512, 0.5, 0, 0, 0, 0)
    h = C.c_void_p()
    rc = lib.t2_ctx_create(C.byref(cfg), C.byref(h))
    assert rc != 0 and b"no CPU fallback" in lib.t2_last_error()
    m = Tacotron2(**synth.VANILLA.kwargs())
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m(torch.ones(1, 3, dtype=torch.long), torch.tensor([3]), False, max_len_override=2)


@pytest.mark.parametrize("args", [synth.VANILLA, synth.HIFI, synth.CONDITIONED])
def test_state_dict_matches_reference_keys(args):
    m = Tacotron2(**args.kwargs())
    sd = synth.synth_state_dict(args)
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert {k: tuple(v.shape) for k, v in m.state_dict().items() if "num_batches" not in k} == \
        {k: tuple(v.shape) for k, v in sd.items() if "num_batches" not in k}
    n = sum(p.numel() for p in m.parameters())
    assert n == {False: 28135729, True: 29332417}[args.description_embeddings] + (2048 if args is synth.HIFI else 0)


def test_signatures_match_reference():
    # model/tacotron2.py:155-166, model/decoder.py:53-67, model/attention.py:34-41
    fwd = list(inspect.signature(Tacotron2.forward).parameters)
    assert fwd[:10] == ["self", "chars_idx", "chars_idx_len", "teacher_forcing", "mel_spectrogram", "mel_spectrogram_len",
                        "speaker_id", "controls", "max_len_override", "description_embeddings"]
    assert list(inspect.signature(Decoder.forward).parameters) == [
        "self", "prev_mel_prenet", "att_rnn_hidden", "att_context", "att_weights", "att_weights_cum", "rnn_hidden", "encoded",
        "att_encoded", "encoded_mask", "speech_features", "extra_att_in", "extra_decoder_in"]
    assert list(inspect.signature(Attention.forward).parameters) == [
        "self", "attention_hidden_state", "memory", "processed_memory", "attention_weights_cat", "mask"]
    assert list(inspect.signature(Encoder.forward).parameters)[:3] == ["self", "char_idx", "char_idx_len"]  # + optional dropout_keep
    assert list(inspect.signature(Postnet.__init__).parameters) == ["self", "num_layers", "num_mels", "postnet_dim", "dropout"]
    d = AlwaysDropout(0.5).eval()
    assert (d(torch.ones(1000)) == 0).any()  # active in eval (model/modules.py:10-12)


def test_reference_asserts():
    m = Tacotron2(**synth.CONDITIONED.kwargs())
    x, n = torch.ones(1, 3, dtype=torch.long), torch.tensor([3])
    with pytest.raises(AssertionError, match="speaker_id tensor required"):
        m(x, n, False, max_len_override=2)
    with pytest.raises(AssertionError, match="Ground-truth Mel spectrogram is required"):
        m(x, n, True)
    v = Tacotron2(**synth.VANILLA.kwargs())
    with pytest.raises(AssertionError, match="Controls are disabled"):
        v(x, n, False, controls=torch.zeros(1, 5), max_len_override=2)
    with pytest.raises(Exception, match="max_len_override is required"):
        v(x, n, False)


def test_environment_switches_are_documented():
    """Every T2B200_* variable the library or the host side reads is listed in INTEGRATION.md (each one selects between two CUDA
    implementations: none of them may appear undocumented, and none may route to a CPU path)."""
    names = set()
    for d in (os.path.join(ROOT, "tacotron2_b200", "csrc"), os.path.join(ROOT, "tacotron2_b200")):
        for f in os.listdir(d):
            if f.endswith((".cu", ".cuh", ".inl", ".py")):
                names |= set(re.findall(r"getenv\(\"(T2B200_[A-Z0-9_]+)\"\)|environ(?:\.get)?[\(\[]\"(T2B200_[A-Z0-9_]+)\"", open(os.path.join(d, f)).read()))
    names = {a or b for a, b in names}
    assert names, "no switches found: the scan is broken"
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    missing = sorted(n for n in names if n not in doc)
    assert not missing, missing
