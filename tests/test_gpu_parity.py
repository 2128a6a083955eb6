"""GPU parity tests: the CUDA path (through the C ABI, via the reference-shaped modules) against
(a) the golden vectors produced by the unmodified reference and (b) the CPU oracle on seeded inputs.

Tolerances are north_star's fp32 bars: max-abs mel <= 1e-3, per-frame attention max-abs <= 1e-4,
identical stop-gate decisions (BASELINE.json).  Most checks pass far tighter bounds; the asserted
value is the contractual one unless a tighter one is stated.
"""
import os

import numpy as np
import pytest
import torch

from oracle import tacotron2_oracle as O
from tacotron2_b200 import synth
from tacotron2_b200.model import Tacotron2

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
MEL_TOL, ATT_TOL = 1e-3, 1e-4


def _dev():
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return torch.device("cuda:0")


def _model(args, sd, train=False):
    m = Tacotron2(**args.kwargs()).to(_dev())
    m.load_state_dict(sd)
    return m.train(train)


def _err(a, b):
    a = a.detach().float().cpu().numpy() if isinstance(a, torch.Tensor) else a
    b = b.detach().float().cpu().numpy() if isinstance(b, torch.Tensor) else b
    assert a.shape == b.shape, (a.shape, b.shape)
    return float(np.max(np.abs(a - b))) if a.size else 0.0


def _cuda(batch):
    return {k: v.to(_dev()) for k, v in batch.items()}


@pytest.mark.parametrize("name,args", [("step_vanilla", synth.VANILLA), ("step_conditioned", synth.CONDITIONED)])
def test_decoder_step_golden(name, args):
    """Decoder.forward (model/decoder.py:53-119) from a non-trivial state, vs the reference's own output."""
    g = dict(np.load(os.path.join(G, name + ".npz")))
    m = _model(args, synth.synth_state_dict(args, seed=0))
    t = lambda k: torch.from_numpy(g["in/" + k]).to(_dev())  # noqa: E731
    w_cum = t("w_cum")
    ctrl = t("controls") if args.controls else None
    with torch.no_grad():
        mel, gate, (ah, ac), ctx, w, wc, (rh, rc) = m.decoder(
            t("prev"), (t("att_h"), t("att_c")), t("ctx"), t("w"), w_cum, (t("rnn_h"), t("rnn_c")), t("memory"), t("pm"),
            t("mask"), extra_decoder_in=ctrl)
    assert wc is w_cum  # in-place update of the caller's tensor (model/decoder.py:90)
    for k, v, tol in (("mel", mel, 1e-5), ("gate", gate, 1e-5), ("att_h", ah, 1e-5), ("att_c", ac, 1e-5), ("ctx", ctx, 1e-5),
                      ("w", w, 1e-6), ("w_cum", wc, 1e-5), ("rnn_h", rh, 1e-5), ("rnn_c", rc, 1e-5)):
        assert _err(v, g[k]) <= tol, (k, _err(v, g[k]))


@pytest.mark.parametrize("name,args", [("ar_vanilla", synth.VANILLA), ("ar_conditioned", synth.CONDITIONED)])
def test_ar_golden(name, args):
    """Tacotron2.forward(teacher_forcing=False) incl. stop logic and output masking vs the reference."""
    g = dict(np.load(os.path.join(G, name + ".npz")))
    B, L, max_len, shift, scale = g["meta"]
    B, L, max_len = int(B), int(L), int(max_len)
    sd = synth.synth_state_dict(args, seed=0, gate_shift=float(shift), gate_scale=float(scale))
    b = _cuda(synth.synth_batch(args, B, L, 8, seed=1234))
    masks = synth.synth_masks(args, B, L, max_len, seed=7)
    m = _model(args, sd)
    with torch.no_grad():
        enc = m.encoder(b["chars_idx"], b["chars_idx_len"])
        mels, post, gates, al = m(b["chars_idx"], b["chars_idx_len"], False, speaker_id=b.get("speaker_id"),
                                  controls=b.get("controls"), max_len_override=max_len,
                                  description_embeddings=b.get("description_embeddings"), dropout_masks=masks)
    assert _err(enc, g["encoded"]) <= 2e-5
    assert mels.shape == g["mels"].shape, "frame count (stop decision) differs"
    assert np.array_equal(gates.cpu().numpy() == -1000.0, g["gates"] == -1000.0), "stop-gate decisions differ"
    assert _err(al, g["alignments"]) <= ATT_TOL
    assert _err(mels, g["mels"]) <= MEL_TOL
    assert _err(post, g["mels_post"]) <= MEL_TOL
    assert _err(al, g["alignments"]) <= 5e-6 and _err(mels, g["mels"]) <= 5e-5  # what fp32 FMA actually achieves


def test_tf_eval_golden():
    """Teacher-forced forward in eval mode (bulk prenet + persistent loop) vs the reference."""
    args = synth.HIFI
    g = dict(np.load(os.path.join(G, "tf_eval_hifi.npz")))
    B, L, T, _ = (int(x) for x in g["meta"])
    b = _cuda(synth.synth_batch(args, B, L, T, seed=1234))
    masks = synth.synth_masks(args, B, L, T, seed=7)
    m = _model(args, synth.synth_state_dict(args, seed=0))
    with torch.no_grad():
        mels, post, gates, al = m(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
                                  speaker_id=b["speaker_id"], dropout_masks=masks)
    assert _err(al, g["alignments"]) <= ATT_TOL
    assert _err(mels, g["mels"]) <= MEL_TOL and _err(post, g["mels_post"]) <= MEL_TOL
    assert _err(gates, g["gates"]) <= MEL_TOL


@pytest.mark.parametrize("B,L", [(1, 1), (2, 5), (7, 64), (64, 200)])
def test_encoder_vs_oracle(B, L):
    """Encoder.forward incl. ragged/unsorted lengths and the packed BiLSTM (model/encoder.py:54-67)."""
    args = synth.VANILLA
    sd = synth.synth_state_dict(args, seed=3)
    b = synth.synth_batch(args, B, L, 4, seed=5)
    m = _model(args, sd)
    with torch.no_grad():
        got = m.encoder(b["chars_idx"].to(_dev()), b["chars_idx_len"].to(_dev()))
    want = O.encoder(sd, b["chars_idx"], b["chars_idx_len"])
    assert _err(got, want) <= 2e-5


@pytest.mark.parametrize("B,T", [(1, 1), (3, 7), (5, 130)])
def test_postnet_vs_oracle(B, T):
    args = synth.VANILLA
    sd = synth.synth_state_dict(args, seed=3)
    x = torch.randn(B, args.num_mels, T, generator=torch.Generator().manual_seed(1))
    m = _model(args, sd)
    with torch.no_grad():
        got = m.postnet(x.to(_dev()))
    want = O.postnet(sd, x)
    assert _err(got, want) <= 5e-5


def test_c2_shape_vs_oracle():
    """BASELINE config C2 shapes (B=64, L=200) for the first 48 frames against the CPU oracle, plus the
    size-independent properties on the full 1000-frame run (the oracle would need minutes there)."""
    args, B, L, T = synth.VANILLA, 64, 200, 48
    sd = synth.synth_state_dict(args, seed=0, gate_shift=5.0)  # never stops
    b = synth.synth_batch(args, B, L, 8, seed=1234)
    masks = synth.synth_masks(args, B, L, 1000, seed=7)
    m = _model(args, sd)
    bc = _cuda(b)
    with torch.no_grad():
        mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks)
    want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], False, max_len_override=T, masks=masks)
    assert _err(al, want[3]) <= ATT_TOL and _err(mels, want[0]) <= MEL_TOL and _err(post, want[1]) <= MEL_TOL
    assert _err(gates, want[2]) <= MEL_TOL
    # full length: properties
    with torch.no_grad():
        mels2, post2, gates2, al2 = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=1000, dropout_masks=masks)
        mels3, _, _, al3 = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=1000, dropout_masks=masks)
    assert mels2.shape == (B, 1000, 80) and al2.shape == (B, 1000, L)
    assert torch.equal(mels2, mels3) and torch.equal(al2, al3), "persistent kernel is not deterministic"
    assert torch.equal(mels2[:, :T], mels), "prefix of the long run differs from the short run"
    assert float((al2.sum(-1) - 1).abs().max()) <= 1e-5  # softmax rows sum to one
    pad = torch.arange(L, device=_dev())[None, :] >= bc["chars_idx_len"][:, None]
    assert float(al2.masked_select(pad[:, None, :].expand_as(al2)).abs().max()) == 0.0  # masked columns exactly 0
    assert torch.isfinite(mels2).all() and torch.isfinite(post2).all()


def test_step_api_matches_persistent_loop():
    """One launch per frame (Decoder.forward + prenet) reproduces the whole-utterance persistent kernel."""
    args, B, L, T = synth.VANILLA, 5, 23, 6
    sd = synth.synth_state_dict(args, seed=0, gate_shift=5.0)
    b = _cuda(synth.synth_batch(args, B, L, 8, seed=11))
    masks = synth.synth_masks(args, B, L, T, seed=7)
    m = _model(args, sd)
    with torch.no_grad():
        mels, _, gates, al = m(b["chars_idx"], b["chars_idx_len"], False, max_len_override=T, dropout_masks=masks)
        enc = m.encoder(b["chars_idx"], b["chars_idx_len"])
        memory, pmem = m._memory(enc, None, None)
        mask = torch.arange(L, device=_dev())[None, :] >= b["chars_idx_len"][:, None]
        ah, ctx, w, wc, rh = m.init_hidden(L, B, _dev())
        prev = torch.zeros(B, args.prenet_dim, device=_dev())
        pk = masks["prenet"].to(_dev()).float() * 2.0
        for t in range(T):
            mel, gate, ah, ctx, w, wc, rh = m.decoder(prev, ah, ctx, w, wc, rh, memory, pmem, mask)
            assert _err(mel, mels[:, t]) <= 1e-5 and _err(w, al[:, t]) <= 1e-6, t
            h1 = torch.relu(mel @ m.prenet[0].weight.T) * pk[t + 1, 0]
            prev = torch.relu(h1 @ m.prenet[3].weight.T) * pk[t + 1, 1]


def test_rng_dropout_statistics():
    """Without injected masks the kernels draw their own AlwaysDropout decisions (p = 0.5)."""
    args = synth.VANILLA
    m = _model(args, synth.synth_state_dict(args, seed=0))
    from tacotron2_b200 import _lib
    lib = m._ready(_dev())
    x = torch.rand(4096, args.num_mels, device=_dev()) + 0.5
    out = torch.empty(4096, args.prenet_dim, device=_dev())
    ws = m._workspace(lib, _dev(), 1, 1, 1)
    _lib.check(lib.t2_prenet_fwd(m._ctx, _lib.ptr(x), None, None, 1234, 4096, _lib.ptr(out), _lib.ptr(ws), ws.numel(), m._stream()))
    ref = torch.relu(torch.relu(x @ m.prenet[0].weight.T) @ m.prenet[3].weight.T)
    frac_zero = float((out == 0).float().mean())
    base_zero = float((ref == 0).float().mean())
    assert 0.45 <= (frac_zero - base_zero) / max(1e-6, 1 - base_zero) <= 0.62


def test_c4_shape_vs_oracle():
    """BASELINE config C4: speaker tokens + 5 controls + 768-d description embeddings (E = 640, K = 896 / 1669 / 1664), batch 32 x 200
    characters, autoregressive: first 24 frames against the CPU oracle, stop decisions on a gate-shifted copy of the weights."""
    args, B, L, T = synth.CONDITIONED, 32, 200, 24
    b = synth.synth_batch(args, B, L, 8, seed=4321)
    bc = _cuda(b)
    masks = synth.synth_masks(args, B, L, T, seed=9)
    for shift, scale in ((5.0, 1.0), (-35.35, -1000.0)):   # never stops / stops at varied frames
        sd = synth.synth_state_dict(args, seed=0, gate_shift=shift, gate_scale=scale)
        m = _model(args, sd)
        kw = dict(speaker_id=bc["speaker_id"], controls=bc["controls"], description_embeddings=bc["description_embeddings"])
        with torch.no_grad():
            mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks, **kw)
        want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], False, speaker_id=b["speaker_id"], controls=b["controls"],
                                   description_embeddings=b["description_embeddings"], max_len_override=T, masks=masks)
        assert mels.shape == want[0].shape, (mels.shape, want[0].shape)   # same number of emitted frames = same stop step
        assert _err(al, want[3]) <= ATT_TOL and _err(mels, want[0]) <= MEL_TOL and _err(post, want[1]) <= MEL_TOL
        assert torch.equal(gates.cpu() == -1000.0, want[2] == -1000.0), "stop decisions differ"


def test_c5_long_form():
    """BASELINE config C5 (long-form stress): batch 16 x 600 characters.  fp32-parity arithmetic: first 12 frames against the CPU
    oracle and the size-independent properties on the full 3000-frame run.  bf16 mode (single tensor-core pass, the mode C5 names):
    stated looser tolerance on the same 12 frames (mel <= 5e-2, attention <= 5e-3) and the same properties at 3000 frames."""
    args, B, L, T, TL = synth.VANILLA, 16, 600, 12, 3000
    sd = synth.synth_state_dict(args, seed=0, gate_shift=5.0)
    b = synth.synth_batch(args, B, L, 8, seed=77)
    bc = _cuda(b)
    masks = synth.synth_masks(args, B, L, TL, seed=3)
    want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], False, max_len_override=T, masks=masks)
    pad = torch.arange(L, device=_dev())[None, :] >= bc["chars_idx_len"][:, None]
    old = os.environ.get("T2B200_PRECISION")
    try:
        for mode, mel_tol, att_tol in (("fp32", MEL_TOL, ATT_TOL), ("bf16", 5e-2, 5e-3)):
            os.environ["T2B200_PRECISION"] = mode
            m = _model(args, sd)
            with torch.no_grad():
                mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks)
                assert _err(al, want[3]) <= att_tol and _err(mels, want[0]) <= mel_tol and _err(post, want[1]) <= mel_tol, mode
                mels2, post2, _, al2 = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=TL, dropout_masks=masks)
                mels3, _, _, al3 = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=TL, dropout_masks=masks)
            assert mels2.shape == (B, TL, 80) and al2.shape == (B, TL, L)
            assert torch.equal(mels2, mels3) and torch.equal(al2, al3), "persistent kernel is not deterministic"
            assert torch.equal(mels2[:, :T], mels), "prefix of the long run differs from the short run"
            assert float((al2.sum(-1) - 1).abs().max()) <= 1e-5
            assert float(al2.masked_select(pad[:, None, :].expand_as(al2)).abs().max()) == 0.0
            assert torch.isfinite(mels2).all() and torch.isfinite(post2).all()
    finally:
        if old is None:
            os.environ.pop("T2B200_PRECISION", None)
        else:
            os.environ["T2B200_PRECISION"] = old


def test_decoder_schedules_agree():
    """Three implementations of the frame loop must give the same results up to summation order, forward and backward: the pipelined
    kernel (decoder_ar.cu: autoregressive synthesis and the training forward; default), decoder_persistent_kernel with the location
    term precomputed beside the GEMMs (T2B200_AR_KERNEL=old), and its in-phase path for very long utterances (+ T2B200_NO_PRECOMPUTE=1).
    The backward consumes three different tapes accordingly (location term taped without / with conv features / not at all)."""
    args, B, L, T = synth.VANILLA, 6, 41, 10
    sd = synth.synth_state_dict(args, seed=0, gate_shift=5.0)
    bc = _cuda(synth.synth_batch(args, B, L, T, seed=5))
    masks_ar = synth.synth_masks(args, B, L, T, seed=7)
    masks_tr = synth.synth_masks(args, B, L, T, seed=7, train=True)
    res = {}
    try:
        for mode, env in (("pipelined", {}), ("persistent", {"T2B200_AR_KERNEL": "old"}),
                          ("persistent-inphase", {"T2B200_AR_KERNEL": "old", "T2B200_NO_PRECOMPUTE": "1"})):
            os.environ.update(env)
            m = _model(args, sd)
            with torch.no_grad():
                ar = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks_ar)
            m.train()
            mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], True, bc["mel_spectrogram"], bc["mel_spectrogram_len"],
                                      dropout_masks=masks_tr)
            m.loss(mels, post, gates, bc["mel_spectrogram"], bc["gate"]).backward()
            res[mode] = (ar, (mels.detach(), al.detach()),
                         {k: p.grad.detach().clone() for k, p in m.named_parameters() if k.startswith(("decoder.", "prenet."))})
            m.check_status()
    finally:
        os.environ.pop("T2B200_NO_PRECOMPUTE", None)
        os.environ.pop("T2B200_AR_KERNEL", None)
    for mode in ("persistent", "persistent-inphase"):
        for a, b_ in zip(res["pipelined"][0], res[mode][0]):
            assert _err(a, b_) <= 2e-5, mode
        for a, b_ in zip(res["pipelined"][1], res[mode][1]):
            assert _err(a, b_) <= 2e-5, mode
        for k, g in res["pipelined"][2].items():
            g2 = res[mode][2][k]
            assert float((g - g2).norm()) <= 2e-4 * float(g.norm()) + 1e-9, (mode, k, float((g - g2).norm()), float(g.norm()))


@pytest.mark.parametrize("with_masks", [True, False])
def test_ar_more_than_64_utterances(with_masks):
    """Autoregressive decode of a batch larger than one persistent-kernel launch (64): the reference's stop condition couples the
    whole batch (model/tacotron2.py:319-322), so every chunk has to produce the frame count of the slowest one.  Stop gates are
    shifted so that the utterances stop at different frames; checked against the CPU oracle run on the whole batch (injected
    masks), and -- with the in-kernel dropout RNG -- against per-chunk consistency properties."""
    args, B, L, T = synth.VANILLA, 70, 19, 24
    sd = synth.synth_state_dict(args, seed=0, gate_shift=-35.35, gate_scale=-1000.0)
    b = synth.synth_batch(args, B, L, 8, seed=1234)
    bc = _cuda(b)
    m = _model(args, sd)
    if with_masks:
        masks = synth.synth_masks(args, B, L, T, seed=7)
        with torch.no_grad():
            mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks)
        want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], False, max_len_override=T, masks=masks)
        assert mels.shape == want[0].shape, ("frame count (stop decision) differs", mels.shape, want[0].shape)
        assert torch.equal(gates.cpu() == -1000.0, want[2] == -1000.0), "stop-gate decisions differ"
        assert _err(al, want[3]) <= ATT_TOL and _err(mels, want[0]) <= MEL_TOL and _err(post, want[1]) <= MEL_TOL
        n_valid = (want[2] != -1000.0).sum(1).ravel()
        assert len(set(n_valid.tolist())) > 1, "test is vacuous: every utterance stopped at the same frame"
    else:
        torch.manual_seed(3)
        with torch.no_grad():
            mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T)
        assert mels.shape[0] == B and mels.shape[1] <= T and torch.isfinite(mels).all() and torch.isfinite(post).all()
        assert float((al.sum(-1) - 1).abs().max()) <= 1e-5


@pytest.mark.parametrize("B,L,T", [(1, 1, 3), (1, 7, 1), (2, 5, 9), (63, 11, 6), (65, 3, 5), (129, 9, 7), (3, 257, 4), (5, 600, 3)])
def test_ar_edge_shapes(B, L, T):
    """Edge shapes of the autoregressive path against the oracle: single utterance / character / frame, batch sizes around the
    64-utterance launch limit (63, 65, 129: chunked decode), encoder lengths beyond one attention tile (257, 600)."""
    args = synth.VANILLA
    sd = synth.synth_state_dict(args, seed=0, gate_shift=-35.35, gate_scale=-1000.0)
    b = synth.synth_batch(args, B, L, 8, seed=1234)
    masks = synth.synth_masks(args, B, L, T, seed=7)
    m = _model(args, sd)
    bc = _cuda(b)
    with torch.no_grad():
        mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], False, max_len_override=T, dropout_masks=masks)
    want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], False, max_len_override=T, masks=masks)
    assert mels.shape == want[0].shape, ("frame count (stop decision) differs", mels.shape, want[0].shape)
    assert torch.equal(gates.cpu() == -1000.0, want[2] == -1000.0), "stop-gate decisions differ"
    assert _err(al, want[3]) <= ATT_TOL and _err(mels, want[0]) <= MEL_TOL and _err(post, want[1]) <= MEL_TOL


@pytest.mark.parametrize("B,L,T", [(1, 1, 1), (65, 7, 5), (3, 300, 4)])
def test_tf_eval_edge_shapes(B, L, T):
    """Teacher-forced eval forward at edge shapes (one frame; more than 64 utterances = two launches; long encoder) vs the oracle."""
    args = synth.VANILLA
    sd = synth.synth_state_dict(args, seed=0)
    b = synth.synth_batch(args, B, L, T, seed=1234)
    masks = synth.synth_masks(args, B, L, T, seed=7)
    m = _model(args, sd)
    bc = _cuda(b)
    with torch.no_grad():
        mels, post, gates, al = m(bc["chars_idx"], bc["chars_idx_len"], True, bc["mel_spectrogram"], bc["mel_spectrogram_len"], dropout_masks=masks)
    want = O.tacotron2_forward(sd, args, b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"], masks=masks)
    assert mels.shape == want[0].shape
    assert _err(al, want[3]) <= ATT_TOL and _err(mels, want[0]) <= MEL_TOL and _err(post, want[1]) <= MEL_TOL
    assert _err(gates, want[2]) <= MEL_TOL


@pytest.mark.parametrize("name,args", [("step_vanilla", synth.VANILLA), ("step_conditioned", synth.CONDITIONED)])
def test_attention_forward_standalone(name, args):
    """Attention.forward called on its own (model/attention.py:34-69, C ABI t2_attention_step): against the reference's step golden
    (the attention inside that Decoder.forward saw query = the new att_h, wcat = (w, w_cum) of the inputs) and against the oracle."""
    g = dict(np.load(os.path.join(G, name + ".npz")))
    sd = synth.synth_state_dict(args, seed=0)
    m = _model(args, sd)
    t = lambda k: torch.from_numpy(g[k]).to(_dev())  # noqa: E731
    wcat = torch.stack([t("in/w"), t("in/w_cum")], 1)
    mask = torch.from_numpy(g["in/mask"]).to(_dev())
    with torch.no_grad():
        ctx, w = m.decoder.attention(t("att_h"), t("in/memory"), t("in/pm"), wcat, mask)
    assert _err(ctx, g["ctx"]) <= 1e-5 and _err(w, g["w"]) <= 1e-6
    assert float(w.masked_select(mask).abs().max() if bool(mask.any()) else 0.0) == 0.0
    # a second shape, against the oracle (bool mask, non-contiguous inputs)
    gen = torch.Generator().manual_seed(4)
    B, L, EF = 7, 301, args.encoded_full_dim
    h = torch.randn(B, args.att_rnn_dim, generator=gen) * 0.3
    mem = torch.randn(B, L, EF, generator=gen) * 0.3
    pm = torch.randn(B, L, args.att_dim, generator=gen) * 0.5
    lens = torch.tensor([L, 250, 17, 1, 300, 128, 64])
    msk = torch.arange(L)[None, :] >= lens[:, None]
    wp = torch.softmax(torch.randn(B, L, generator=gen).masked_fill(msk, float("-inf")), 1)
    wc = wp * 3.0
    want_ctx, want_w = O.attention(sd, h, mem, pm, wp, wc, msk)
    with torch.no_grad():
        ctx, w = m.decoder.attention(h.to(_dev()), mem.to(_dev()), pm.to(_dev()), torch.stack([wp, wc], 1).to(_dev()), msk.to(_dev()))
    assert _err(ctx, want_ctx) <= 1e-5 and _err(w, want_w) <= 1e-6
