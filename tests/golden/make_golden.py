"""Generate golden vectors from the UNMODIFIED reference (run in the build container only).

    python tests/golden/make_golden.py          # needs /root/reference

Imports ``model.tacotron2.Tacotron2`` from /root/reference, loads the seeded
weights of ``tacotron2_b200.synth`` into it, injects the seeded dropout masks by
swapping ``nn.Dropout`` slots on the *instance* (reference source untouched,
SURVEY.md section 8c) and stores the outputs as small ``.npz`` files next to this
script.  Inputs/weights/masks are NOT stored: the tests regenerate them from the
same seeds.  /root/reference does not exist on the GPU box, so nothing but this
script reads it.
"""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
sys.path.insert(0, "/root/reference")

from tacotron2_b200 import synth  # noqa: E402

with contextlib.redirect_stdout(io.StringIO()):
    from model.tacotron2 import Tacotron2  # noqa: E402  (the reference)


class Stream(torch.nn.Module):
    """Stands in for nn.Dropout / AlwaysDropout: multiplies by the next injected mask."""

    def __init__(self, masks, p, always):
        super().__init__()
        self.masks, self.p, self.always, self.i = masks, p, always, 0

    def forward(self, x):
        if not (self.always or self.training):
            return x
        m = self.masks[self.i]
        self.i += 1
        return x * m.to(x.dtype) * (1.0 / (1.0 - self.p))


def build(args, sd):
    with contextlib.redirect_stdout(io.StringIO()):
        m = Tacotron2(**args.kwargs())
    missing = m.load_state_dict(sd, strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return m


def inject(m, args, masks, teacher_forcing, train):
    pk = masks["prenet"]  # (T+1,2,B,P)
    if teacher_forcing:  # one bulk call on (B,T+1,80)  model/tacotron2.py:255-257
        m.prenet[2] = Stream([pk[:, 0].permute(1, 0, 2)], args.dropout, True)
        m.prenet[5] = Stream([pk[:, 1].permute(1, 0, 2)], args.dropout, True)
    else:  # one call per frame  model/tacotron2.py:263,325
        m.prenet[2] = Stream(list(pk[:, 0]), args.dropout, True)
        m.prenet[5] = Stream(list(pk[:, 1]), args.dropout, True)
    if train:
        m.decoder.att_rnn_dropout = Stream(list(masks["att_h"]), 0.1, False)
        m.decoder.lstm_dropout = Stream(list(masks["rnn_h"]), 0.1, False)
        for li, slot in enumerate((3, 7, 11)):
            m.encoder.convolutions[slot] = Stream([masks["encoder"][li].transpose(1, 2)], args.dropout, False)
        for li, slot in enumerate((3, 7, 11, 15, 18)):
            c = args.num_mels if li == 4 else args.postnet_dim
            m.postnet.postnet[slot] = Stream([masks["postnet"][li][:, :, :c].transpose(1, 2)], args.dropout, False)


def np_(t):
    return t.detach().cpu().numpy()


def case_ar(name, args, B, L, max_len, gate_shift, gate_scale):
    sd = synth.synth_state_dict(args, seed=0, gate_shift=gate_shift, gate_scale=gate_scale)
    batch = synth.synth_batch(args, B, L, 8, seed=1234)
    masks = synth.synth_masks(args, B, L, max_len, seed=7)
    m = build(args, sd).eval()
    inject(m, args, masks, False, False)
    with torch.no_grad():
        enc = m.encoder(batch["chars_idx"], batch["chars_idx_len"])
        mels, post, gates, al = m(
            batch["chars_idx"], batch["chars_idx_len"], False,
            speaker_id=batch.get("speaker_id"), controls=batch.get("controls"),
            max_len_override=max_len, description_embeddings=batch.get("description_embeddings"),
        )
    print(name, "frames", mels.shape[1], "valid lengths", (gates[:, :, 0] > -999).sum(1).tolist())
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), encoded=np_(enc), mels=np_(mels), mels_post=np_(post), gates=np_(gates),
        alignments=np_(al), meta=np.array([B, L, max_len, gate_shift, gate_scale], np.float64),
    )


def case_tf(name, args, B, L, T, train):
    sd = synth.synth_state_dict(args, seed=0)
    batch = synth.synth_batch(args, B, L, T, seed=1234)
    masks = synth.synth_masks(args, B, L, T, seed=7, train=train)
    m = build(args, sd)
    m.train(train)
    inject(m, args, masks, True, train)
    mels, post, gates, al = m(
        batch["chars_idx"], batch["chars_idx_len"], True, batch["mel_spectrogram"], batch["mel_spectrogram_len"],
        speaker_id=batch.get("speaker_id"), controls=batch.get("controls"),
        description_embeddings=batch.get("description_embeddings"),
    )
    F = torch.nn.functional  # model/tts_model.py:197-201
    loss = (
        F.binary_cross_entropy_with_logits(gates, batch["gate"])
        + F.mse_loss(mels, batch["mel_spectrogram"])
        + F.mse_loss(post, batch["mel_spectrogram"])
    )
    out = dict(mels=np_(mels), mels_post=np_(post), gates=np_(gates), alignments=np_(al), loss=np_(loss),
               meta=np.array([B, L, T, int(train)], np.float64))
    if train:
        loss.backward()
        for k, p in m.named_parameters():
            g = p.grad if p.grad is not None else torch.zeros_like(p)
            g = g.reshape(-1)
            # summary: [sum, sum|g|, sum g^2] + a strided sample of 64 entries
            idx = torch.linspace(0, g.numel() - 1, min(64, g.numel())).long()
            out["grad/" + k] = np.concatenate([np_(torch.stack([g.sum(), g.abs().sum(), (g * g).sum()])), np_(g[idx])])
        # BN running stats after one train step (momentum 0.1, unbiased var)
        for k, v in m.state_dict().items():
            if "running_" in k:
                out["bn/" + k] = np_(v)
    print(name, "loss", float(loss.detach()))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)


def case_step(name, args, B, L):
    """One Decoder.forward from a non-trivial state (model/decoder.py:53-119)."""
    sd = synth.synth_state_dict(args, seed=0)
    m = build(args, sd).eval()
    g = torch.Generator().manual_seed(99)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    EF, H, R = args.encoded_full_dim, args.att_rnn_dim, args.rnn_hidden_dim
    lens = torch.tensor([L - 3 * b for b in range(B)])
    mask = torch.arange(L)[None, :] >= lens[:, None]
    w = torch.softmax(r(B, L).masked_fill(mask, float("-inf")), 1)
    w_cum = w * 2.5
    ins = dict(prev=r(B, args.prenet_dim).relu(), att_h=r(B, H) * 0.3, att_c=r(B, H) * 0.5, ctx=r(B, EF) * 0.2, w=w,
               w_cum=w_cum.clone(), rnn_h=r(B, R) * 0.3, rnn_c=r(B, R) * 0.5, memory=r(B, L, EF) * 0.3,
               pm=r(B, L, args.att_dim) * 0.5, mask=mask)
    ctrl = r(B, args.controls_dim).clamp(-1, 1) if args.controls else None
    with torch.no_grad():
        mel, gate, (ah, ac), ctx, w2, wc2, (rh, rc) = m.decoder(
            ins["prev"], (ins["att_h"], ins["att_c"]), ins["ctx"], ins["w"], ins["w_cum"].clone(),
            (ins["rnn_h"], ins["rnn_c"]), ins["memory"], ins["pm"], ins["mask"], extra_decoder_in=ctrl)
    out = {"in/" + k: np_(v) for k, v in ins.items()}
    if ctrl is not None:
        out["in/controls"] = np_(ctrl)
    out.update(mel=np_(mel), gate=np_(gate), att_h=np_(ah), att_c=np_(ac), ctx=np_(ctx), w=np_(w2), w_cum=np_(wc2),
               rnn_h=np_(rh), rnn_c=np_(rc))
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, "ok")


if __name__ == "__main__":
    torch.manual_seed(0)
    torch.set_num_threads(8)
    case_step("step_vanilla", synth.VANILLA, 5, 33)
    case_step("step_conditioned", synth.CONDITIONED, 3, 19)
    case_ar("ar_vanilla", synth.VANILLA, 4, 37, 40, -35.35, -1000.0)
    case_ar("ar_conditioned", synth.CONDITIONED, 3, 21, 12, 0.0, 1.0)
    case_tf("tf_eval_hifi", synth.HIFI, 3, 26, 14, False)
    case_tf("tf_train_vanilla", synth.VANILLA, 3, 29, 17, True)
    case_tf("tf_train_conditioned", synth.CONDITIONED, 2, 18, 9, True)
    case_tf("tf_train_long", synth.VANILLA, 4, 40, 200, True)
