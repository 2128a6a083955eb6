"""t2_trim_static: the allocation-free variant of t2_trim (scratch in a file-scope device array of T2_TRIM_STATIC_MAX_FRAMES floats,
include/t2b200.h).  No counterpart in the reference -- so the checker is the existing
entry point t2_trim on the same inputs; the outputs are int64 sample bounds, so the bar is equality.  Two calls in a row on different
batches (the second smaller and shorter, so that the shared array still holds the first call's frames), the size limit, and that the
torch allocator is not touched.  The file name sorts last on purpose.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from tacotron2_b200 import _lib, synth

pytestmark = pytest.mark.gpu
MAX_FRAMES = 262144   # T2_TRIM_STATIC_MAX_FRAMES


def _batch(specs, dev):
    wavs = []
    for i, (n, lead, tail, seed) in enumerate(specs):
        w = synth.synth_wav(n, seed=seed)
        w = np.concatenate([1e-6 * np.sin(np.arange(lead)), w, np.zeros(tail)]).astype(np.float32)
        wavs.append(w)
    stride = max(len(w) for w in wavs)
    pad = torch.zeros(len(wavs), stride)
    for i, w in enumerate(wavs):
        pad[i, : len(w)] = torch.from_numpy(w)
    return pad.to(dev), torch.tensor([len(w) for w in wavs], dtype=torch.int64, device=dev), stride


def _existing(lib, pad, lens, stride, top_db, fl, hop):
    B = pad.shape[0]
    ws = torch.empty(lib.t2_trim_workspace_bytes(B, stride, hop), dtype=torch.uint8, device=pad.device)
    s, e = torch.empty(B, dtype=torch.int64, device=pad.device), torch.empty(B, dtype=torch.int64, device=pad.device)
    _lib.check(lib.t2_trim(_lib.ptr(pad), _lib.ptr(lens), B, stride, stride, top_db, fl, hop, _lib.ptr(s), _lib.ptr(e), _lib.ptr(ws), ws.numel(),
                           C.c_void_p(torch.cuda.current_stream().cuda_stream)), "t2_trim")
    return s, e


def _static(lib, pad, lens, stride, top_db, fl, hop, s, e):
    return lib.t2_trim_static(_lib.ptr(pad), _lib.ptr(lens), pad.shape[0], stride, stride, top_db, fl, hop, _lib.ptr(s), _lib.ptr(e),
                              C.c_void_p(torch.cuda.current_stream().cuda_stream))


def test_trim_static_matches_trim_on_two_calls_in_a_row():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    first = [(22050, 3000, 4000, 0), (9000, 0, 2500, 1), (5120, 700, 0, 2), (2048, 100, 100, 3), (40000, 12345, 777, 4), (30000, 5000, 9000, 5)]
    second = [(7000, 1500, 600, 11), (12000, 0, 0, 12), (3000, 2048, 4096, 13)]
    with torch.cuda.device(dev):
        for specs, top_db, fl in ((first, 60.0, 2048), (second, 30.0, 1024)):
            pad, lens, stride = _batch(specs, dev)
            want_s, want_e = _existing(lib, pad, lens, stride, top_db, fl, 512)
            s = torch.full((len(specs),), -7, dtype=torch.int64, device=dev)
            e = torch.full((len(specs),), -7, dtype=torch.int64, device=dev)
            torch.cuda.synchronize()
            before = (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev), torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0))
            _lib.check(_static(lib, pad, lens, stride, top_db, fl, 512, s, e), "t2_trim_static")
            torch.cuda.synchronize()
            after = (torch.cuda.memory_allocated(dev), torch.cuda.memory_reserved(dev), torch.cuda.memory_stats(dev).get("allocation.all.allocated", 0))
            assert before == after, ("the call went through torch's allocator", before, after)
            assert torch.equal(s, want_s) and torch.equal(e, want_e), (s.tolist(), want_s.tolist(), e.tolist(), want_e.tolist())
            assert bool((e > s).any()), "degenerate fixture: nothing kept"
        # the same inputs once more: the array now holds exactly this call's frames
        s2, e2 = torch.empty_like(s), torch.empty_like(e)
        _lib.check(_static(lib, pad, lens, stride, top_db, fl, 512, s2, e2), "t2_trim_static")
        assert torch.equal(s2, want_s) and torch.equal(e2, want_e)


def test_trim_static_rejects_sizes_beyond_the_built_in_scratch():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    with torch.cuda.device(dev):
        pad, lens, stride = _batch([(4096, 0, 0, 0)], dev)
        s = torch.full((1,), -7, dtype=torch.int64, device=dev)
        e = torch.full((1,), -7, dtype=torch.int64, device=dev)
        # B * (1 + max_len / hop) just over the limit: rejected on the host before any launch (no buffer is dereferenced)
        B_big = 64
        max_len = (MAX_FRAMES // B_big) * 512          # 1 + max_len / 512 = MAX_FRAMES / 64 + 1 frames per utterance
        rc = lib.t2_trim_static(_lib.ptr(pad), _lib.ptr(lens), B_big, max_len, max_len, 60.0, 2048, 512, _lib.ptr(s), _lib.ptr(e),
                                C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == -4, rc   # T2_ERR_WORKSPACE
        msg = lib.t2_last_error().decode()
        assert "t2_trim_static" in msg and str(MAX_FRAMES) in msg, msg
        torch.cuda.synchronize()
        assert s.item() == -7 and e.item() == -7, "a rejected call must not launch"
        # exactly at the limit is accepted: one utterance whose frame count is MAX_FRAMES (only `lens` samples are ever read)
        max_len = (MAX_FRAMES - 1) * 512
        big = torch.zeros(1, max_len, device=dev)
        big[0, 100000:200000] = torch.from_numpy(synth.synth_wav(100000, seed=5)).to(dev)
        blen = torch.tensor([max_len], dtype=torch.int64, device=dev)
        want_s, want_e = _existing(lib, big, blen, max_len, 60.0, 2048, 512)
        _lib.check(lib.t2_trim_static(_lib.ptr(big), _lib.ptr(blen), 1, max_len, max_len, 60.0, 2048, 512, _lib.ptr(s), _lib.ptr(e),
                                      C.c_void_p(torch.cuda.current_stream().cuda_stream)), "t2_trim_static")
        assert torch.equal(s, want_s) and torch.equal(e, want_e) and e.item() > s.item() > 0
