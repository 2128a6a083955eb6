"""Lightning-free mirror of the reference's training wrapper.

``TTSModel`` keeps the constructor keywords, ``forward`` signature and the ``training_step`` / ``validation_step`` /
``predict_step`` / ``configure_optimizers`` protocol of ``model/tts_model.py:19-115,116-202,304-327`` so that a maintainer can
drive it from ``run/train.py`` without ``lightning`` (absent in this image): the batch is the reference's collate output
``(tts_data, tts_metadata, tts_extra)`` (``datasets/tts_dataloader.py:8-35``).  The optimizer shell is the fused
``ClipAdam`` (Adam + L2 + global-norm clip 1.0, ``run/train.py:239-240``) with the reference's ``MultiStepLR(gamma=0.1)``
stepped per training step (``model/tts_model.py:85-90``).  ``from_config`` follows ``run/train.py:176-228``;
``load_reference_checkpoint`` takes a Lightning ``.ckpt`` (keys prefixed ``tacotron2.``) or a bare ``Tacotron2`` state dict.
"""
from __future__ import annotations

import json
from typing import Iterable, List, Optional

import torch
from torch import Tensor, nn

from .model import Tacotron2
from .optim import ClipAdam


class MultiStepLR:
    """``torch.optim.lr_scheduler.MultiStepLR`` arithmetic for :class:`ClipAdam`: lr = base * gamma ** #{milestones <= step}."""

    def __init__(self, optimizer, milestones: Iterable[int], gamma: float = 0.1, base_lr: Optional[float] = None):
        self.optimizer, self.milestones, self.gamma = optimizer, sorted(int(m) for m in milestones), float(gamma)
        self.base_lr = float(optimizer.lr if base_lr is None else base_lr)
        self.last_epoch = 0

    def lr_at(self, step: int) -> float:
        return self.base_lr * self.gamma ** sum(1 for m in self.milestones if m <= step)

    def step(self) -> None:
        self.last_epoch += 1
        self.optimizer.set_lr(self.lr_at(self.last_epoch))

    def state_dict(self):
        from collections import Counter
        # torch.optim.lr_scheduler.MultiStepLR.state_dict() keys (milestones as a Counter, base_lrs, _last_lr) + this class's own
        return {"last_epoch": self.last_epoch, "base_lr": self.base_lr, "milestones": Counter(self.milestones), "gamma": self.gamma,
                "base_lrs": [self.base_lr], "_last_lr": [self.lr_at(self.last_epoch)], "_step_count": self.last_epoch + 1}

    def load_state_dict(self, sd):
        self.last_epoch = int(sd["last_epoch"])
        self.base_lr = float(sd["base_lr"] if "base_lr" in sd else sd["base_lrs"][0])
        ms = sd["milestones"]
        self.milestones = sorted(int(m) for m in (ms.elements() if hasattr(ms, "elements") else ms))
        self.gamma = float(sd["gamma"])
        self.optimizer.set_lr(self.lr_at(self.last_epoch))


class TTSModel(nn.Module):
    def __init__(self, lr: float, weight_decay: float, num_chars: int, encoded_dim: int, encoder_kernel_size: int, num_mels: int,
                 prenet_dim: int, att_rnn_dim: int, att_dim: int, rnn_hidden_dim: int, postnet_dim: int, dropout: float,
                 scheduler_milestones: List[int], speaker_tokens: bool = False, num_speakers: int = 1, controls: bool = False,
                 controls_dim: int = 0, max_len_override: Optional[int] = None, description_embeddings: bool = False,
                 description_embeddings_dim: int = 0, gradient_clip_val: float = 1.0):
        super().__init__()
        self._hparams = {k: v for k, v in locals().items() if k not in ("self", "__class__", "gradient_clip_val")}   # save_hyperparameters()
        self.lr, self.weight_decay, self.scheduler_milestones = lr, weight_decay, list(scheduler_milestones)
        self.speaker_tokens, self.controls, self.max_len_override = speaker_tokens, controls, max_len_override
        self.description_embeddings = description_embeddings
        self.gradient_clip_val = gradient_clip_val
        self.global_step = 0
        self.tacotron2 = Tacotron2(num_chars=num_chars, encoded_dim=encoded_dim, encoder_kernel_size=encoder_kernel_size,
                                   num_mels=num_mels, prenet_dim=prenet_dim, att_rnn_dim=att_rnn_dim, att_dim=att_dim,
                                   rnn_hidden_dim=rnn_hidden_dim, postnet_dim=postnet_dim, dropout=dropout,
                                   speaker_tokens=speaker_tokens, num_speakers=num_speakers, controls=controls,
                                   controls_dim=controls_dim, description_embeddings=description_embeddings,
                                   description_embeddings_dim=description_embeddings_dim)

    # ------------------------------------------------------------------ construction from the reference's files
    @classmethod
    def from_config(cls, config, **overrides) -> "TTSModel":
        """``config``: a dict or the path of one of the reference's ``config/*.json`` (``run/train.py:176-228``)."""
        if not isinstance(config, dict):
            with open(config) as f:
                config = json.load(f)
        ds, tr, mc = config["dataset"]["preprocessing"], config["training"], config["model"]
        ext = config.get("extensions", {})
        ctl, spk = ext.get("controls", {}), ext.get("speaker_tokens", {})   # description embeddings come through model.args
        args = dict(mc["args"])
        if "char_embedding_dim" in args:   # the shipped configs still carry the older keyword for encoded_dim
            args["encoded_dim"] = args.pop("char_embedding_dim")
        max_steps = tr.get("args", {}).get("max_steps", 0)
        kw = dict(lr=tr["lr"], weight_decay=tr["weight_decay"],
                  num_chars=len(ds["allowed_chars"]) + (ds.get("end_token") is not None), num_mels=ds["num_mels"],
                  controls=bool(ctl.get("active")), controls_dim=len(ctl.get("features", [])) if ctl.get("active") else 0,
                  speaker_tokens=bool(spk.get("active")), num_speakers=spk.get("num_speakers", 1) if spk.get("active") else 1,
                  scheduler_milestones=[int(x * max_steps) for x in mc.get("scheduler_milestones", [])], **args)
        kw.update(overrides)
        return cls(**kw)

    def load_reference_checkpoint(self, ckpt, strict: bool = True):
        """A Lightning checkpoint (path or dict with ``state_dict``, keys ``tacotron2.*``) or a bare ``Tacotron2`` state dict."""
        if not isinstance(ckpt, dict):
            ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
        sd = ckpt.get("state_dict", ckpt)
        if any(k.startswith("tacotron2.") for k in sd):
            sd = {k[len("tacotron2."):]: v for k, v in sd.items() if k.startswith("tacotron2.")}
        out = self.tacotron2.load_state_dict(sd, strict=strict)
        if "global_step" in ckpt:
            self.global_step = int(ckpt["global_step"])
        return out

    def save_checkpoint(self, path, optimizer=None, scheduler=None, epoch: int = 0):
        """What ``trainer.save_checkpoint`` leaves behind (run/train.py:245-255), as far as this model determines it: a dict with
        ``state_dict`` under the LightningModule's ``tacotron2.`` prefix (the keys ``TTSModel.load_from_checkpoint`` of
        run/say.py:125-137 reads), ``global_step`` / ``epoch``, ``hyper_parameters`` (the constructor keywords of
        model/tts_model.py:19-43), and the optimizer / scheduler state in torch's own layouts under Lightning's keys.
        ``load_reference_checkpoint`` + ``resume`` read it back."""
        self.tacotron2.check_status()   # never write a checkpoint after an aborted kernel
        sd = {"tacotron2." + k: v.detach().cpu().clone() for k, v in self.tacotron2.state_dict().items()}
        ckpt = {"epoch": int(epoch), "global_step": int(self.global_step), "pytorch-lightning_version": "2.0.0", "state_dict": sd,
                "loops": {}, "callbacks": {}, "hparams_name": "kwargs", "hyper_parameters": dict(self._hparams),
                "optimizer_states": [], "lr_schedulers": []}
        if optimizer is not None:
            osd = optimizer.state_dict()
            osd["state"] = {i: {k: (v.detach().cpu().clone() if torch.is_tensor(v) else v) for k, v in st.items()}
                            for i, st in osd["state"].items()}
            ckpt["optimizer_states"] = [osd]
        if scheduler is not None:
            ckpt["lr_schedulers"] = [scheduler.state_dict()]
        torch.save(ckpt, path)
        return ckpt

    def resume(self, ckpt, optimizer=None, scheduler=None):
        """``Trainer.fit(ckpt_path=...)`` (run/train.py:245-249): weights, step counter, optimizer moments, scheduler position."""
        if not isinstance(ckpt, dict):
            ckpt = torch.load(ckpt, map_location="cpu", weights_only=False)
        self.load_reference_checkpoint(ckpt)
        if optimizer is not None and ckpt.get("optimizer_states"):
            optimizer.load_state_dict(ckpt["optimizer_states"][0])
        if scheduler is not None and ckpt.get("lr_schedulers"):
            scheduler.load_state_dict(ckpt["lr_schedulers"][0])
        return ckpt

    # ------------------------------------------------------------------ reference protocol
    def forward(self, chars_idx: Tensor, chars_idx_len: Tensor, teacher_forcing: bool = True, mel_spectrogram: Optional[Tensor] = None,
                mel_spectrogram_len: Optional[Tensor] = None, speaker_id: Optional[Tensor] = None, controls: Optional[Tensor] = None,
                max_len_override: Optional[int] = None, description_embeddings: Optional[Tensor] = None):
        return self.tacotron2(chars_idx=chars_idx, chars_idx_len=chars_idx_len, teacher_forcing=teacher_forcing,
                              mel_spectrogram=mel_spectrogram, mel_spectrogram_len=mel_spectrogram_len, speaker_id=speaker_id,
                              controls=controls, max_len_override=max_len_override, description_embeddings=description_embeddings)

    def configure_optimizers(self):
        opt = ClipAdam(self.tacotron2.parameters(), lr=self.lr, weight_decay=self.weight_decay, max_norm=self.gradient_clip_val)
        cfg = {"optimizer": opt}
        if len(self.scheduler_milestones) > 0:
            cfg["lr_scheduler"] = {"scheduler": MultiStepLR(opt, self.scheduler_milestones, gamma=0.1), "interval": "step"}
        return cfg

    def _conditioning(self, tts_metadata):
        args = {}
        if self.speaker_tokens:
            args["speaker_id"] = tts_metadata["speaker_id"]
        if self.controls:
            args["controls"] = tts_metadata["features"]
        if self.description_embeddings:
            args["description_embeddings"] = tts_metadata["description_embeddings"]
        return args

    def _teacher_forced(self, batch):
        tts_data, tts_metadata, _ = batch
        mel, post, gate, alignment = self(chars_idx=tts_data["chars_idx"], chars_idx_len=tts_metadata["chars_idx_len"], teacher_forcing=True,
                                          mel_spectrogram=tts_data["mel_spectrogram"], mel_spectrogram_len=tts_metadata["mel_spectrogram_len"],
                                          **self._conditioning(tts_metadata))
        # gate BCE-with-logits + mel MSE + post-net mel MSE, plain means (model/tts_model.py:197-201) as one fused kernel
        loss = self.tacotron2.loss(mel, post, gate, tts_data["mel_spectrogram"], tts_data["gate"])
        return loss, mel, post, gate, alignment

    def training_step(self, batch, batch_idx=0):
        loss, *_ = self._teacher_forced(batch)
        return loss

    def validation_step(self, batch, batch_idx=0):
        tts_data, tts_metadata, _ = batch
        with torch.no_grad():
            loss, mel, post, gate, alignment = self._teacher_forced(batch)
        n, m = int(tts_metadata["mel_spectrogram_len"][0]), int(tts_metadata["chars_idx_len"][0])
        return {"mel_spectrogram_pred": post[0, :n], "mel_spectrogram": tts_data["mel_spectrogram"][0, :n], "alignment": alignment[0, :n, :m],
                "gate": tts_data["gate"][0], "gate_pred": gate[0], "loss": loss}

    def predict_step(self, batch, batch_idx=0, dataloader_idx=None):
        tts_data, tts_metadata, tts_extra = batch
        args = self._conditioning(tts_metadata)
        args.pop("description_embeddings", None)   # the reference's predict_step does not pass them (model/tts_model.py:304-327)
        with torch.no_grad():
            mel, post, gate, alignment = self(chars_idx=tts_data["chars_idx"], chars_idx_len=tts_metadata["chars_idx_len"], teacher_forcing=False,
                                              mel_spectrogram=tts_data["mel_spectrogram"], mel_spectrogram_len=tts_metadata["mel_spectrogram_len"],
                                              max_len_override=self.max_len_override, **args)
        return mel, post, gate, alignment, tts_extra.get("text")


def fit(model: TTSModel, batches, max_steps: int, grad_sync=None, on_step=None):
    """What ``Trainer.fit`` does for this model, without the framework: ``training_step`` -> ``backward`` -> (rank-averaged
    gradients when ``grad_sync`` is a :class:`tacotron2_b200.parallel.GradSync`) -> clip + Adam -> scheduler, until
    ``max_steps``; ``batches`` is any iterable of collated batches already on the device (re-iterated when exhausted)."""
    cfg = model.configure_optimizers()
    opt, sched = cfg["optimizer"], cfg.get("lr_scheduler", {}).get("scheduler")
    model.train()
    while model.global_step < max_steps:
        seen = False
        for i, batch in enumerate(batches):
            seen = True
            opt.zero_grad()
            loss = model.training_step(batch, i)
            loss.backward()
            if grad_sync is not None:
                grad_sync.finish()
            norm = opt.step()
            if sched is not None:
                sched.step()
            model.global_step += 1
            if on_step is not None:
                on_step(model.global_step, loss, norm)
                model.tacotron2.check_status()   # on_step reads the loss (a host sync anyway): an aborted kernel raises here
            if model.global_step >= max_steps:
                break
        if not seen:
            raise ValueError("fit: empty batch iterable")
    return opt, sched
