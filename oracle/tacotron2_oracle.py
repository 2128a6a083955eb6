"""CPU oracle for the Tacotron2 acoustic-model hot path.  TEST INFRASTRUCTURE ONLY.

This file restates, function by function, what mattm458/tacotron2 computes on
the path Encoder -> autoregressive Decoder (prenet, attention LSTMCell,
location-sensitive attention, decoder LSTMCell, mel/gate projection) -> Postnet
and the training loss, as plain fp32 torch-on-CPU arithmetic with every dropout
mask injected explicitly.  Only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it; the
product (``tacotron2_b200``) never does.

Parity status: PINNED against the reference itself.  ``tests/golden/make_golden.py``
imports the unmodified reference modules from /root/reference in the build
container, feeds them the seeded weights/batches/masks of
``tacotron2_b200.synth`` and commits their outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every function here against those files.
(The reference has no tests or golden vectors of its own)

Parameters are passed as a flat dict with the reference's state-dict keys
(SURVEY.md section 8b).  Masks are uint8 keep-masks in the layouts of
``tacotron2_b200.synth.synth_masks``; ``None`` means "dropout inactive".
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

BN_EPS = 1e-5  # nn.BatchNorm1d default (model/encoder.py:41, model/postnet.py:16)
HID_P = 0.1  # hard-coded nn.Dropout(0.1) on att_h / rnn_h (model/decoder.py:29,43)


def _drop(x, keep, p):
    """nn.Dropout / AlwaysDropout with an injected keep-mask: x * keep / (1-p).
    (model/modules.py:10-12; torch dropout scales kept values by 1/(1-p).)"""
    if keep is None:
        return x
    return x * keep.to(x.dtype) * (1.0 / (1.0 - p))


# ----------------------------------------------------------------------------- prenet
def prenet(sd, x, keep1, keep2, p):
    """model/tacotron2.py:85-92: Linear(no bias) -> ReLU -> AlwaysDropout, twice."""
    h = _drop(torch.relu(x @ sd["prenet.0.weight"].T), keep1, p)
    return _drop(torch.relu(h @ sd["prenet.3.weight"].T), keep2, p)


# ----------------------------------------------------------------------------- LSTM cell
def lstm_cell(x, h, c, w_ih, w_hh, b_ih, b_hh):
    """nn.LSTMCell arithmetic (model/decoder.py:26-28,40-42): gate order i,f,g,o."""
    gates = x @ w_ih.T + b_ih + h @ w_hh.T + b_hh
    i, f, g, o = gates.chunk(4, dim=1)
    c2 = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
    h2 = torch.sigmoid(o) * torch.tanh(c2)
    return h2, c2


# ----------------------------------------------------------------------------- attention
def attention(sd, att_h, memory, processed_memory, w_prev, w_cum, mask, pre="decoder.attention."):
    """model/attention.py:34-69.  Channel 0 = previous weights, channel 1 =
    cumulative weights (model/decoder.py:78-80).  Returns (context, weights)."""
    q = att_h @ sd[pre + "query_layer.weight"].T  # (B,A)                      :52
    wcat = torch.stack([w_prev, w_cum], dim=1)  # (B,2,L)
    loc = F.conv1d(wcat, sd[pre + "location_conv.weight"], padding=15)  # (B,32,L)   :54
    loc = loc.transpose(1, 2) @ sd[pre + "location_dense.weight"].T  # (B,L,A)       :55-56
    e = torch.tanh(q[:, None, :] + loc + processed_memory) @ sd[pre + "v.weight"][0]  # (B,L)  :58-62
    e = e.masked_fill(mask, float("-inf"))  #                                         :63
    w = torch.softmax(e, dim=1)  #                                                    :65
    ctx = torch.bmm(w[:, None, :], memory)[:, 0]  #                                   :66-67
    return ctx, w


# ----------------------------------------------------------------------------- decoder step
def decoder_step(sd, x_prenet, state, memory, processed_memory, mask, controls=None, keep_att=None, keep_rnn=None):
    """One frame, model/decoder.py:53-119.  ``state`` = dict(att_h, att_c, ctx,
    w, w_cum, rnn_h, rnn_c); returns (mel, gate, new_state).  Does not mutate."""
    p = "decoder."
    att_h, att_c = lstm_cell(
        torch.cat([x_prenet, state["ctx"]], -1), state["att_h"], state["att_c"],
        sd[p + "att_rnn.weight_ih"], sd[p + "att_rnn.weight_hh"], sd[p + "att_rnn.bias_ih"], sd[p + "att_rnn.bias_hh"],
    )  # :70-74
    att_h = _drop(att_h, keep_att, HID_P)  # :75 (dropped h is carried and used as query)
    ctx, w = attention(sd, att_h, memory, processed_memory, state["w"], state["w_cum"], mask)  # :78-87
    w_cum = state["w_cum"] + w  # :90
    dec_in = [att_h, ctx] + ([controls] if controls is not None else [])  # :94-98
    rnn_h, rnn_c = lstm_cell(
        torch.cat(dec_in, -1), state["rnn_h"], state["rnn_c"],
        sd[p + "lstm.weight_ih"], sd[p + "lstm.weight_hh"], sd[p + "lstm.bias_ih"], sd[p + "lstm.bias_hh"],
    )  # :100
    rnn_h = _drop(rnn_h, keep_rnn, HID_P)  # :101
    hc = torch.cat([rnn_h, ctx], -1)
    gate = hc @ sd[p + "gate.weight"].T + sd[p + "gate.bias"]  # :103-104 (controls NOT in gate)
    mel_in = torch.cat([hc, controls], -1) if controls is not None else hc
    mel = mel_in @ sd[p + "mel_out.weight"].T + sd[p + "mel_out.bias"]  # :106-109
    new = dict(att_h=att_h, att_c=att_c, ctx=ctx, w=w, w_cum=w_cum, rnn_h=rnn_h, rnn_c=rnn_c)
    return mel, gate, new


def init_state(B, L, H, R, EF):
    """model/tacotron2.py:126-153: all zeros."""
    z = torch.zeros
    return dict(att_h=z(B, H), att_c=z(B, H), ctx=z(B, EF), w=z(B, L), w_cum=z(B, L), rnn_h=z(B, R), rnn_c=z(B, R))


# ----------------------------------------------------------------------------- batch norm
def batchnorm(x, sd, pre, train):
    """nn.BatchNorm1d over (B,C,T) (model/encoder.py:41, model/postnet.py:16).
    train=True: biased batch statistics over (B,T) INCLUDING padded positions
    (SURVEY appendix 13).  Returns (y, batch_mean, batch_var_biased)."""
    w, b = sd[pre + "weight"], sd[pre + "bias"]
    if train:
        mean = x.mean(dim=(0, 2))
        var = x.var(dim=(0, 2), unbiased=False)
    else:
        mean, var = sd[pre + "running_mean"], sd[pre + "running_var"]
    y = (x - mean[None, :, None]) * torch.rsqrt(var[None, :, None] + BN_EPS) * w[None, :, None] + b[None, :, None]
    return y, mean, var


# ----------------------------------------------------------------------------- encoder
def encoder_convs(sd, chars_idx, train=False, keep=None, p=0.5):
    """model/encoder.py:55-57: embedding -> 3x[Conv1d k5 pad2 (bias) -> BN -> ReLU -> Dropout].
    ``keep``: (3,B,L,E) channels-last keep-masks or None.  Returns (B,L,E)."""
    x = F.embedding(chars_idx, sd["encoder.embedding.weight"], padding_idx=0).transpose(1, 2)  # (B,E,L); row 0 gets no grad (:25)
    for li, i in enumerate((0, 4, 8)):
        k = sd[f"encoder.convolutions.{i}.weight"].shape[2]
        x = F.conv1d(x, sd[f"encoder.convolutions.{i}.weight"], sd[f"encoder.convolutions.{i}.bias"], padding=int((k - 1) / 2))
        x, _, _ = batchnorm(x, sd, f"encoder.convolutions.{i + 1}.", train)
        x = torch.relu(x)
        x = _drop(x, None if keep is None else keep[li].transpose(1, 2), p)
    return x.transpose(1, 2)


def bilstm_packed(sd, x, lens, pre="encoder.lstm."):
    """model/encoder.py:59-65: BiLSTM on a packed sequence.  Forward direction runs
    t = 0..len-1, reverse runs t = len-1..0 from a zero state at each utterance's
    own last character; outputs at t >= len are zero (pad_packed_sequence)."""
    B, L, _ = x.shape
    Hh = sd[pre + "weight_hh_l0"].shape[1]
    out = torch.zeros(B, L, 2 * Hh)
    for d, sfx in enumerate(("", "_reverse")):
        w_ih, w_hh = sd[pre + "weight_ih_l0" + sfx], sd[pre + "weight_hh_l0" + sfx]
        b_ih, b_hh = sd[pre + "bias_ih_l0" + sfx], sd[pre + "bias_hh_l0" + sfx]
        h = torch.zeros(B, Hh)
        c = torch.zeros(B, Hh)
        steps = range(L) if d == 0 else range(L - 1, -1, -1)
        for t in steps:
            act = (t < lens)[:, None]
            h2, c2 = lstm_cell(x[:, t], h, c, w_ih, w_hh, b_ih, b_hh)
            h = torch.where(act, h2, h)
            c = torch.where(act, c2, c)
            out[:, t, d * Hh:(d + 1) * Hh] = torch.where(act, h2, torch.zeros_like(h2))
    return out


def encoder(sd, chars_idx, chars_idx_len, train=False, keep=None, p=0.5):
    """model/encoder.py:54-67."""
    conv = encoder_convs(sd, chars_idx, train, keep, p)
    out = bilstm_packed(sd, conv, chars_idx_len)
    return out[:, : int(chars_idx_len.max())]  # pad_packed_sequence pads to the longest length


# ----------------------------------------------------------------------------- postnet
def postnet(sd, x, train=False, keep=None, p=0.5):
    """model/postnet.py:8-52 on (B,M,T): 5x Conv1d k5 'same' no bias -> BN -> (Tanh on first 4) -> Dropout.
    ``keep``: (5,B,T,512) channels-last keep-masks (layer 5 uses the first M channels)."""
    for li, i in enumerate((0, 4, 8, 12, 16)):
        x = F.conv1d(x, sd[f"postnet.postnet.{i}.weight"], None, padding=2)
        x, _, _ = batchnorm(x, sd, f"postnet.postnet.{i + 1}.", train)
        if li < 4:
            x = torch.tanh(x)
        if keep is not None:
            x = _drop(x, keep[li][:, :, : x.shape[1]].transpose(1, 2), p)
    return x


# ----------------------------------------------------------------------------- conditioning
def condition(sd, encoded, speaker_id=None, description=None):
    """model/tacotron2.py:201-212: tanh(enc + speaker_emb) over ALL positions, then
    concat Tanh(Linear(description)) broadcast over L."""
    if speaker_id is not None:
        encoded = torch.tanh(encoded + sd["speaker_embedding.weight"][speaker_id][:, None, :])
    if description is not None:
        d = torch.tanh(description @ sd["description_embeddings_linear.0.weight"].T + sd["description_embeddings_linear.0.bias"])
        encoded = torch.cat([encoded, d[:, None, :].expand(-1, encoded.shape[1], -1)], dim=2)
    return encoded


# ----------------------------------------------------------------------------- full forward
def tacotron2_forward(
    sd, args, chars_idx, chars_idx_len, teacher_forcing, mel_spectrogram=None, mel_spectrogram_len=None,
    speaker_id=None, controls=None, max_len_override=None, description_embeddings=None,
    masks=None, train=False, return_states=False,
):
    """model/tacotron2.py:155-347.  ``args`` is a ``synth.ModelArgs``; ``train`` selects
    BatchNorm batch statistics and the train-only dropouts (their masks must then be
    in ``masks``).  The prenet masks are ALWAYS required (AlwaysDropout).
    Returns (mels, mels_post, gates, alignments[, per-step states])."""
    masks = masks or {}
    B, L = chars_idx.shape
    p = args.dropout
    encoded = encoder(sd, chars_idx, chars_idx_len, train, masks.get("encoder") if train else None, p)  # :197
    encoded = condition(
        sd, encoded,
        speaker_id if args.speaker_tokens else None,
        description_embeddings if args.description_embeddings else None,
    )  # :201-212
    mask = torch.arange(L)[None, :] >= chars_idx_len[:, None]  # :223-226
    att_encoded = encoded @ sd["att_encoder.weight"].T  # :229
    state = init_state(B, encoded.shape[1], args.att_rnn_dim, args.rnn_hidden_dim, args.encoded_full_dim)  # :240
    if max_len_override is not None:
        max_len = max_len_override
    elif mel_spectrogram is not None:
        max_len = mel_spectrogram.shape[1]
    else:
        raise Exception("If Mel spectrogram is not given, max_len_override is required!")
    pk = masks["prenet"]  # (T+1,2,B,P)
    if teacher_forcing:
        dec_in = F.pad(mel_spectrogram, (0, 0, 1, 0))  # :255
        # one bulk draw in the reference (B,T+1,P); injected here per frame -- same values
        dec_in = torch.stack([prenet(sd, dec_in[:, t], pk[t, 0], pk[t, 1], p) for t in range(dec_in.shape[1])], 1)
        prev = dec_in[:, 0]
        lengths = mel_spectrogram_len
    else:
        prev = prenet(sd, torch.zeros(B, args.num_mels), pk[0, 0], pk[0, 1], p)  # :262-263
        done = torch.zeros(B, dtype=torch.bool)
        lengths = torch.zeros(B, dtype=torch.long)
    mels, gates, aligns, states = [], [], [], []
    for i in range(max_len):  # :276
        ka = masks["att_h"][i] if train else None
        kr = masks["rnn_h"][i] if train else None
        mel, gate, state = decoder_step(sd, prev, state, encoded, att_encoded, mask, controls if args.controls else None, ka, kr)
        mels.append(mel)
        gates.append(gate)
        aligns.append(state["w"])
        if return_states:
            states.append({k: v.detach().clone() for k, v in state.items()})
        if teacher_forcing:
            prev = dec_in[:, i + 1]  # :317
        else:
            g = gate[:, 0]
            done = done | (g < 0.0)  # :319
            lengths = lengths + (g >= 0.0).long()  # :320 (keeps counting after done)
            if bool(done.all()):  # :321-322 (the breaking frame is emitted)
                break
            prev = prenet(sd, mel.detach(), pk[i + 1, 0], pk[i + 1, 1], p)  # :324-325
    mels = torch.stack(mels, 1)
    gates = torch.stack(gates, 1)
    aligns = torch.stack(aligns, 1)
    post = postnet(sd, mels.transpose(1, 2), train, masks.get("postnet") if train else None, p).transpose(1, 2)  # :332
    mels_post = mels + post  # :333 (unmasked mels)
    mel_mask = torch.arange(mels.shape[1])[None, :] >= lengths[:, None]  # :335-338
    mels = mels.masked_fill(mel_mask[:, :, None], 0.0)  # :343
    mels_post = mels_post.masked_fill(mel_mask[:, :, None], 0.0)  # :344
    gates = gates.masked_fill(mel_mask[:, :, None], -1000.0)  # :345
    out = (mels, mels_post, gates, aligns)
    return out + (states,) if return_states else out


def loss_fn(mels, mels_post, gates, mel_target, gate_target):
    """model/tts_model.py:197-201: plain means over the padded tensors."""
    return (
        F.binary_cross_entropy_with_logits(gates, gate_target)
        + F.mse_loss(mels, mel_target)
        + F.mse_loss(mels_post, mel_target)
    )
