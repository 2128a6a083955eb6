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


def _config_dir():
    import os
    for d in ("/root/reference/config", os.path.join(os.path.dirname(__file__), "..", "oracle", "_ref", "config")):
        if os.path.isdir(d):
            return d
    return None


def test_every_reference_config_maps():
    """Every config/*.json the reference ships goes through from_config (run/train.py:176-228), including the drift SURVEY section 5
    lists: the older `char_embedding_dim` keyword, `extensions.prosody_model` / `extensions.descriptions` blocks the HEAD code
    ignores, `description_embeddings` passed through model.args, speaker_tokens.dim.  Parameter shapes are checked against the
    reference's own constructor when it is staged."""
    import glob
    import json
    import os
    d = _config_dir()
    if d is None:
        pytest.skip("reference configs not available (neither /root/reference nor oracle/_ref)")
    files = sorted(glob.glob(os.path.join(d, "*.json")))
    assert len(files) >= 16
    from oracle import ref_model as R
    seen = 0
    for f in files:
        cfg = json.load(open(f))
        if "model" not in cfg:      # server.json: the web demo's deployment table, not a model config
            continue
        m = TTSModel.from_config(f)
        seen += 1
        ext, ds = cfg["extensions"], cfg["dataset"]["preprocessing"]
        sd = m.tacotron2.state_dict()
        assert sd["encoder.embedding.weight"].shape[0] == len(ds["allowed_chars"]) + (ds["end_token"] is not None) + 1, f
        assert ("speaker_embedding.weight" in sd) == bool(ext["speaker_tokens"]["active"]), f
        if ext["speaker_tokens"]["active"]:
            assert sd["speaker_embedding.weight"].shape[0] == ext["speaker_tokens"]["num_speakers"], f
        C = len(ext["controls"]["features"]) if ext["controls"]["active"] else 0
        EF = 512 + (128 if cfg["model"]["args"].get("description_embeddings") else 0)
        assert sd["decoder.lstm.weight_ih"].shape == (4096, 1024 + EF + C), f
        assert m.scheduler_milestones == [int(x * cfg["training"]["args"]["max_steps"]) for x in cfg["model"]["scheduler_milestones"]], f
        if R.available():
            args = dict(cfg["model"]["args"])
            if "char_embedding_dim" in args:
                args["encoded_dim"] = args.pop("char_embedding_dim")
            ref = R._import("model.tacotron2").Tacotron2(
                num_chars=len(ds["allowed_chars"]) + (ds["end_token"] is not None), num_mels=ds["num_mels"], controls=C > 0, controls_dim=C,
                speaker_tokens=bool(ext["speaker_tokens"]["active"]),
                num_speakers=ext["speaker_tokens"].get("num_speakers", 1) if ext["speaker_tokens"]["active"] else 1, **args)
            rsd = ref.state_dict()
            assert list(rsd.keys()) == list(sd.keys()), f
            assert all(rsd[k].shape == sd[k].shape for k in sd), f
    assert seen >= 15


def test_save_checkpoint_roundtrip_cpu(tmp_path):
    """save_checkpoint writes the Lightning-shaped dict run/say.py:125-137 loads: `tacotron2.`-prefixed state_dict, global_step,
    hyper_parameters; load_reference_checkpoint and (when staged) the reference's own Tacotron2 accept it."""
    m = TTSModel.from_config(CONFIG)
    m.load_reference_checkpoint(synth.synth_state_dict(synth.VANILLA, seed=5))
    m.global_step = 77
    p = str(tmp_path / "final.ckpt")
    m.save_checkpoint(p)
    ck = torch.load(p, map_location="cpu", weights_only=False)
    assert ck["global_step"] == 77 and all(k.startswith("tacotron2.") for k in ck["state_dict"])
    assert ck["hyper_parameters"]["num_chars"] == 39 and ck["hyper_parameters"]["scheduler_milestones"] == [500, 750]
    m2 = TTSModel(**ck["hyper_parameters"])     # what load_from_checkpoint does with the saved hyper-parameters
    m2.load_reference_checkpoint(p)
    assert m2.global_step == 77
    for k, v in m.tacotron2.state_dict().items():
        assert torch.equal(v, m2.tacotron2.state_dict()[k]), k
    from oracle import ref_model as R
    if R.available():
        ref = R._import("model.tacotron2").Tacotron2(**synth.VANILLA.kwargs())
        ref.load_state_dict({k[len("tacotron2."):]: v for k, v in ck["state_dict"].items()}, strict=True)


def test_deepcopy_and_pickle_cpu():
    import copy
    import pickle
    m = TTSModel.from_config(CONFIG)
    m2 = copy.deepcopy(m)
    assert m2.tacotron2.encoder._t2_owner() is m2.tacotron2 and m.tacotron2.encoder._t2_owner() is m.tacotron2
    assert torch.equal(m2.tacotron2.prenet[0].weight, m.tacotron2.prenet[0].weight)
    m3 = pickle.loads(pickle.dumps(m.tacotron2))
    assert m3.decoder.attention._t2_owner() is m3 and m3._ctx is None


@pytest.mark.gpu
def test_resume_continues_the_run(tmp_path):
    """2 steps + save + resume + 2 steps == 4 steps (weights, Adam moments, step counter, scheduler position, BatchNorm counters)."""
    dev = torch.device("cuda:0")
    c = copy.deepcopy(CONFIG)
    c["training"]["args"]["max_steps"] = 4
    b = {k: v.to(dev) for k, v in synth.synth_batch(synth.VANILLA, 4, 23, 17, seed=5).items()}
    batch = ({"chars_idx": b["chars_idx"], "mel_spectrogram": b["mel_spectrogram"], "gate": b["gate"]},
             {"chars_idx_len": b["chars_idx_len"], "mel_spectrogram_len": b["mel_spectrogram_len"]}, {"text": ["x"] * 4})

    def fresh():
        m = TTSModel.from_config(c).to(dev)
        m.load_reference_checkpoint(synth.synth_state_dict(synth.VANILLA, seed=0))
        return m

    torch.manual_seed(11)
    a = fresh()
    fit(a, [batch], max_steps=4)
    torch.manual_seed(11)
    h = fresh()
    opt, sched = fit(h, [batch], max_steps=2)
    rng = torch.get_rng_state()
    p = str(tmp_path / "mid.ckpt")
    h.save_checkpoint(p, opt, sched)
    r = fresh()
    cfg = r.configure_optimizers()
    r.resume(p, cfg["optimizer"], cfg["lr_scheduler"]["scheduler"])
    assert r.global_step == 2 and cfg["optimizer"].step_count == 2 and abs(cfg["optimizer"].lr - 1e-4) <= 1e-12
    torch.set_rng_state(rng)
    # continue with the restored optimizer / scheduler (fit() would build new ones)
    r.train()
    for i in range(2):
        cfg["optimizer"].zero_grad()
        r.training_step(batch, i).backward()
        cfg["optimizer"].step()
        cfg["lr_scheduler"]["scheduler"].step()
        r.global_step += 1
    sa, sr = a.tacotron2.state_dict(), r.tacotron2.state_dict()
    for k in sa:
        assert torch.equal(sa[k], sr[k]), k
    assert int(sa["encoder.convolutions.1.num_batches_tracked"]) == 4   # counted like nn.BatchNorm1d does
    assert deepcopy_runs(r)


def deepcopy_runs(m):
    """A deep copy of a model that has a live CUDA context gets its own context lazily and computes the same thing."""
    dev = torch.device("cuda:0")
    m2 = copy.deepcopy(m).eval()
    m.eval()
    b = {k: v.to(dev) for k, v in synth.synth_batch(synth.VANILLA, 2, 9, 4, seed=1).items()}
    masks = synth.synth_masks(synth.VANILLA, 2, 9, 5, seed=2)
    with torch.no_grad():
        x = m.tacotron2(b["chars_idx"], b["chars_idx_len"], False, max_len_override=5, dropout_masks=masks)[1]
        y = m2.tacotron2(b["chars_idx"], b["chars_idx_len"], False, max_len_override=5, dropout_masks=masks)[1]
    return torch.equal(x, y)
