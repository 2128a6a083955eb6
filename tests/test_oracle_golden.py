"""Pins the CPU oracle against outputs of the unmodified reference (tests/golden/*.npz,
written by tests/golden/make_golden.py from /root/reference)."""
import os

import numpy as np
import pytest
import torch

from oracle import tacotron2_oracle as O
from tacotron2_b200 import synth

G = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return dict(np.load(os.path.join(G, name + ".npz")))


def _close(a, b, tol, what):
    a = a.detach().numpy() if isinstance(a, torch.Tensor) else a
    assert a.shape == b.shape, (what, a.shape, b.shape)
    err = float(np.max(np.abs(a - b))) if a.size else 0.0
    assert err <= tol, f"{what}: max-abs {err} > {tol}"


@pytest.mark.parametrize("name,args", [("step_vanilla", synth.VANILLA), ("step_conditioned", synth.CONDITIONED)])
def test_decoder_step(name, args):
    g = _load(name)
    sd = synth.synth_state_dict(args, seed=0)
    t = lambda k: torch.from_numpy(g["in/" + k])  # noqa: E731
    st = dict(att_h=t("att_h"), att_c=t("att_c"), ctx=t("ctx"), w=t("w"), w_cum=t("w_cum"), rnn_h=t("rnn_h"), rnn_c=t("rnn_c"))
    ctrl = t("controls") if args.controls else None
    mel, gate, new = O.decoder_step(sd, t("prev"), st, t("memory"), t("pm"), t("mask"), ctrl)
    _close(mel, g["mel"], 2e-6, "mel")
    _close(gate, g["gate"], 2e-6, "gate")
    for k in ("att_h", "att_c", "ctx", "w", "w_cum", "rnn_h", "rnn_c"):
        _close(new[k], g[k], 2e-6, k)


@pytest.mark.parametrize("name,args", [("ar_vanilla", synth.VANILLA), ("ar_conditioned", synth.CONDITIONED)])
def test_ar(name, args):
    g = _load(name)
    B, L, max_len, shift, scale = g["meta"]
    B, L, max_len = int(B), int(L), int(max_len)
    sd = synth.synth_state_dict(args, seed=0, gate_shift=float(shift), gate_scale=float(scale))
    b = synth.synth_batch(args, B, L, 8, seed=1234)
    masks = synth.synth_masks(args, B, L, max_len, seed=7)
    enc = O.encoder(sd, b["chars_idx"], b["chars_idx_len"])
    _close(enc, g["encoded"], 5e-6, "encoded")
    mels, post, gates, al = O.tacotron2_forward(
        sd, args, b["chars_idx"], b["chars_idx_len"], False, speaker_id=b.get("speaker_id"), controls=b.get("controls"),
        max_len_override=max_len, description_embeddings=b.get("description_embeddings"), masks=masks)
    _close(mels, g["mels"], 1e-5, "mels")
    _close(post, g["mels_post"], 2e-5, "mels_post")
    _close(al, g["alignments"], 1e-6, "alignments")
    # identical stop decisions: the -1000 fill pattern and the frame count
    assert np.array_equal(gates.numpy() == -1000.0, g["gates"] == -1000.0)
    _close(gates, g["gates"], 2e-3 if abs(scale) > 1 else 1e-5, "gates")


@pytest.mark.parametrize(
    "name,args,train",
    [("tf_eval_hifi", synth.HIFI, False), ("tf_train_vanilla", synth.VANILLA, True), ("tf_train_conditioned", synth.CONDITIONED, True)],
)
def test_tf(name, args, train):
    g = _load(name)
    B, L, T, _ = (int(x) for x in g["meta"])
    sd = synth.synth_state_dict(args, seed=0)
    b = synth.synth_batch(args, B, L, T, seed=1234)
    masks = synth.synth_masks(args, B, L, T, seed=7, train=train)
    if train:
        sd = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v) for k, v in sd.items()}
    mels, post, gates, al = O.tacotron2_forward(
        sd, args, b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"],
        speaker_id=b.get("speaker_id"), controls=b.get("controls"), description_embeddings=b.get("description_embeddings"),
        masks=masks, train=train)
    _close(mels, g["mels"], 2e-5, "mels")
    _close(post, g["mels_post"], 1e-4, "mels_post")
    _close(gates, g["gates"], 2e-5, "gates")
    _close(al, g["alignments"], 1e-6, "alignments")
    loss = O.loss_fn(mels, post, gates, b["mel_spectrogram"], b["gate"])
    assert abs(float(loss.detach()) - float(g["loss"])) <= 1e-4 * abs(float(g["loss"]))
    if train:
        loss.backward()
        for k in g:
            if not k.startswith("grad/"):
                continue
            p = sd[k[5:]]
            gr = (p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
            idx = torch.linspace(0, gr.numel() - 1, min(64, gr.numel())).long()
            ref = g[k]
            scale = max(float(np.sqrt(ref[2])), 1e-8)  # ||g||_2 of the reference gradient
            assert abs(float(gr.sum()) - ref[0]) <= 2e-3 * scale + 1e-6, k
            assert abs(float((gr * gr).sum()) - ref[2]) <= 1e-3 * ref[2] + 1e-10, k
            assert float(np.max(np.abs(gr[idx].numpy() - ref[3:]))) <= 1e-3 * scale + 1e-7, k
