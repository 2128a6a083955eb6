"""Host logic of the Lightning-free TTSModel mirror (model/tts_model.py, run/train.py:176-228): config mapping, checkpoint key
mapping, MultiStepLR arithmetic (CPU); one short fit() on the GPU."""
import copy

import pytest
import torch

from tacotron2_b200 import synth
from tacotron2_b200.tts_model import MultiStepLR, TTSModel, fit

CONFIG = {   # shape of the reference's config/*.json (values of vanilla-ljspeech-stop, smaller max_steps)
    "dataset": {"preprocessing": {"allowed_chars": "!'(),.:;? \\-abcdefghijklmnopqrstuvwxyz", "end_token": "^", "num_mels": 80}},
    "training": {"lr": 0.001, "batch_size": 32, "weight_decay": 0.000001, "precision": "16-mixed", "args": {"max_steps": 1000}},
    "model": {"scheduler_milestones": [0.5, 0.75],
              "args": {"prenet_dim": 256, "att_rnn_dim": 1024, "att_dim": 128, "rnn_hidden_dim": 1024, "postnet_dim": 512, "dropout": 0.5,
                       "char_embedding_dim": 512, "encoder_kernel_size": 5}},
    "extensions": {"speaker_tokens": {"active": False}, "controls": {"active": False}},
}


def test_from_config_vanilla():
    m = TTSModel.from_config(CONFIG)
    assert m.scheduler_milestones == [500, 750] and m.lr == 0.001 and m.weight_decay == 1e-6
    sd = m.tacotron2.state_dict()
    assert sd["encoder.embedding.weight"].shape == (39 + 1, 512)     # 38 allowed chars + end token, + padding row
    assert sd["decoder.att_rnn.weight_ih"].shape == (4096, 768) and "speaker_embedding.weight" not in sd


def test_from_config_extensions():
    c = copy.deepcopy(CONFIG)
    c["extensions"] = {"speaker_tokens": {"active": True, "num_speakers": 4}, "controls": {"active": True, "features": list("abcde")}}
    c["model"]["args"].update(description_embeddings=True, description_embeddings_dim=768)
    c["model"]["args"]["encoded_dim"] = c["model"]["args"].pop("char_embedding_dim")
    m = TTSModel.from_config(c)
    sd = m.tacotron2.state_dict()
    assert sd["speaker_embedding.weight"].shape == (4, 512)
    assert sd["decoder.lstm.weight_ih"].shape == (4096, 1024 + 640 + 5) and sd["decoder.gate.weight"].shape == (1, 1024 + 640)
    assert m.speaker_tokens and m.controls and m.description_embeddings


def test_load_reference_checkpoint_prefix():
    a = synth.VANILLA
    m = TTSModel.from_config(CONFIG)
    sd = synth.synth_state_dict(a, seed=3)
    ckpt = {"state_dict": {"tacotron2." + k: v for k, v in sd.items()}, "global_step": 1234}
    m.load_reference_checkpoint(ckpt)
    assert m.global_step == 1234
    assert torch.equal(m.tacotron2.state_dict()["decoder.lstm.weight_hh"], sd["decoder.lstm.weight_hh"])
    m.load_reference_checkpoint(sd)   # bare Tacotron2 state dict


def test_multistep_lr_matches_torch():
    class Opt:
        lr = 1e-3

        def set_lr(self, v):
            self.lr = v

    o = Opt()
    s = MultiStepLR(o, [3, 6], gamma=0.1)
    p = torch.nn.Parameter(torch.zeros(1))
    ref_o = torch.optim.SGD([p], lr=1e-3)
    ref = torch.optim.lr_scheduler.MultiStepLR(ref_o, milestones=[3, 6], gamma=0.1)
    for _ in range(9):
        ref_o.step(); ref.step(); s.step()  # noqa: E702
        assert abs(o.lr - ref_o.param_groups[0]["lr"]) <= 1e-12
    s2 = MultiStepLR(Opt(), [1])
    s2.load_state_dict(s.state_dict())
    assert s2.last_epoch == 9 and abs(s2.optimizer.lr - o.lr) <= 1e-15


@pytest.mark.gpu
def test_fit_runs_reference_protocol():
    dev = torch.device("cuda:0")
    c = copy.deepcopy(CONFIG)
    c["training"]["args"]["max_steps"] = 4
    m = TTSModel.from_config(c).to(dev)
    m.load_reference_checkpoint(synth.synth_state_dict(synth.VANILLA, seed=0))
    b = {k: v.to(dev) for k, v in synth.synth_batch(synth.VANILLA, 4, 23, 17, seed=5).items()}
    batch = ({"chars_idx": b["chars_idx"], "mel_spectrogram": b["mel_spectrogram"], "gate": b["gate"]},
             {"chars_idx_len": b["chars_idx_len"], "mel_spectrogram_len": b["mel_spectrogram_len"]}, {"text": ["x"] * 4})
    seen = []
    opt, sched = fit(m, [batch], max_steps=4, on_step=lambda step, loss, norm: seen.append((step, float(loss), float(norm))))
    assert m.global_step == 4 and [s[0] for s in seen] == [1, 2, 3, 4]
    assert all(torch.isfinite(torch.tensor(s[1:])).all() for s in seen) and seen[-1][1] < seen[0][1]
    assert abs(opt.lr - 1e-3 * 0.1 * 0.1) <= 1e-12          # milestones int(0.5*4)=2 and int(0.75*4)=3 both passed
    out = m.validation_step(batch)
    assert out["alignment"].shape == (int(b["mel_spectrogram_len"][0]), int(b["chars_idx_len"][0]))
    m.eval()
    m.max_len_override = 9
    mel, post, gate, al, text = m.predict_step(batch)
    assert post.shape[0] == 4 and post.shape[1] <= 9 and text == ["x"] * 4
