"""Loader for the UNMODIFIED reference staged under ``oracle/_ref/`` (see ``oracle/stage_ref.py``).  TEST / BASELINE
INFRASTRUCTURE ONLY: ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s reference legs may import this module; the
product never does.

``reference_tacotron2(args, sd)`` builds ``model/tacotron2.py:Tacotron2`` with the seeded weights of ``tacotron2_b200.synth``;
``inject_masks`` swaps the ``nn.Dropout`` / ``AlwaysDropout`` slots of the *instance* for mask streams (reference source
untouched) so that the reference, the oracle and the CUDA path see identical dropout draws.
"""
from __future__ import annotations

import contextlib
import importlib
import io
import os
import sys

import torch

REF_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def available() -> bool:
    return os.path.exists(os.path.join(REF_ROOT, "model", "tacotron2.py"))


def _import(name):
    """Import ``name`` (e.g. ``model.tacotron2``) from oracle/_ref without leaving oracle/_ref on sys.path."""
    if not available():
        raise RuntimeError("oracle/_ref is not staged: run `python oracle/stage_ref.py` where /root/reference exists")
    sys.path.insert(0, REF_ROOT)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            return importlib.import_module(name)
    finally:
        sys.path.remove(REF_ROOT)


class MaskStream(torch.nn.Module):
    """Stands in for nn.Dropout / AlwaysDropout: multiplies by the next injected keep-mask, scaled by 1/(1-p)."""

    def __init__(self, masks, p, always):
        super().__init__()
        self.masks, self.p, self.always, self.i = masks, p, always, 0

    def forward(self, x):
        if not (self.always or self.training):
            return x
        m = self.masks[self.i]
        self.i += 1
        return x * m.to(device=x.device, dtype=x.dtype) * (1.0 / (1.0 - self.p))


def reference_tacotron2(args, sd, device="cpu"):
    mod = _import("model.tacotron2")
    with contextlib.redirect_stdout(io.StringIO()):
        m = mod.Tacotron2(**args.kwargs())
    res = m.load_state_dict(sd, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    return m.to(device)


def inject_masks(m, args, masks, teacher_forcing, train):
    """Same slots as tests/golden/make_golden.py (model/tacotron2.py:85-92, model/decoder.py:29,43, model/encoder.py:43,
    model/postnet.py:18,32,45)."""
    pk = masks["prenet"]  # (T+1,2,B,P)
    if teacher_forcing:  # one bulk call on (B,T+1,80)  model/tacotron2.py:255-257
        m.prenet[2] = MaskStream([pk[:, 0].permute(1, 0, 2)], args.dropout, True)
        m.prenet[5] = MaskStream([pk[:, 1].permute(1, 0, 2)], args.dropout, True)
    else:  # one call per frame  model/tacotron2.py:263,325
        m.prenet[2] = MaskStream(list(pk[:, 0]), args.dropout, True)
        m.prenet[5] = MaskStream(list(pk[:, 1]), args.dropout, True)
    if train:
        m.decoder.att_rnn_dropout = MaskStream(list(masks["att_h"]), 0.1, False)
        m.decoder.lstm_dropout = MaskStream(list(masks["rnn_h"]), 0.1, False)
        for li, slot in enumerate((3, 7, 11)):
            m.encoder.convolutions[slot] = MaskStream([masks["encoder"][li].transpose(1, 2)], args.dropout, False)
        for li, slot in enumerate((3, 7, 11, 15, 18)):
            c = args.num_mels if li == 4 else args.postnet_dim
            m.postnet.postnet[slot] = MaskStream([masks["postnet"][li][:, :, :c].transpose(1, 2)], args.dropout, False)
    return m


def reference_loss(mels, mels_post, gates, mel_target, gate_target):
    """model/tts_model.py:197-201 (that file needs `lightning`, absent here: its three lines are restated)."""
    F = torch.nn.functional
    return F.binary_cross_entropy_with_logits(gates, gate_target) + F.mse_loss(mels, mel_target) + F.mse_loss(mels_post, mel_target)


def reference_generator(h):
    """model/hifi_gan.py:154-216 Generator(h) -- h: attribute dict of the HiFi-GAN config (run/say.py:76-90)."""
    return _import("model.hifi_gan").Generator(h)


def reference_collate():
    return _import("datasets.tts_dataloader")._collate


def config_dir():
    return os.path.join(REF_ROOT, "config")
