"""GPU batch assembly behind the reference's collate interface (datasets/tts_dataloader.py:8-35).

``collate(data, device)`` takes what the reference's ``_collate`` takes -- a list of ``(data, lengths, extra)`` dict triples as
``TTSDataset.__getitem__`` returns them (datasets/tts_dataset.py:226-302) -- and returns the same ``(tts_data, tts_data_len,
tts_data_extra)`` triple, with the padded batch built on the GPU by one ``t2_collate`` launch: each field's items are concatenated
into one pinned host buffer, copied once, and padded on the device (rows beyond an item's length are zero, exactly
``pad_sequence(batch_first=True)``).  ``gate`` may be omitted from the items: it is then generated on the device from the
mel lengths (ones, last valid frame 0: datasets/tts_dataset.py:211-214).  No CPU fallback: without the CUDA library this raises.
"""
from __future__ import annotations

import ctypes as C
from collections import defaultdict

import torch

from . import _lib

MAX_FIELDS = 8


def _table(ctype, vals):
    return (ctype * len(vals))(*vals)


class CollatePlan:
    """One prepared ``t2_collate`` launch: the concatenated fields and offset tables are on the device, the outputs are allocated."""

    def __init__(self, fields: dict, device, gate_from: str | None = None, gate_len_dev: torch.Tensor | None = None, gate_T: int = 0):
        self.lib = _lib.load()
        device = torch.device(device)
        if device.type != "cuda":
            raise RuntimeError("tacotron2_b200.collate: CUDA device required (no CPU fallback)")
        names = list(fields)
        if len(names) > MAX_FIELDS:
            raise ValueError(f"at most {MAX_FIELDS} fields per launch")
        self.device, self.B = device, (len(fields[names[0]]) if names else 0)
        B = self.B
        self.keep, src, off, rowb, maxr, outp, self.outs = [], [], [], [], [], [], {}
        self.bytes_read = self.bytes_written = 0
        gate_len = None
        for k in names:
            items = fields[k]
            assert len(items) == B, f"field {k}: {len(items)} items, expected {B}"
            v0 = items[0]
            trailing = tuple(v0.shape[1:])
            row_elems = 1
            for t in trailing:
                row_elems *= int(t)
            row_bytes = v0.element_size() * row_elems
            if row_bytes % 4:
                raise ValueError(f"field {k}: row size {row_bytes} B is not a multiple of 4")
            lens = [int(v.shape[0]) for v in items]
            o = torch.zeros(B + 1, dtype=torch.int64)
            o[1:] = torch.tensor(lens).cumsum(0)
            cat = torch.cat([v.reshape(v.shape[0], -1) for v in items]).contiguous()
            if cat.device.type != "cuda":
                cat = cat.pin_memory().to(device, non_blocking=True)
            od = o.to(device, non_blocking=True)
            out = torch.empty((B, max(lens)) + trailing, dtype=v0.dtype, device=device)
            self.keep += [cat, od]
            src.append(_lib.ptr(cat)); off.append(_lib.ptr(od)); rowb.append(row_bytes); maxr.append(max(lens)); outp.append(_lib.ptr(out))
            self.outs[k] = out
            self.bytes_read += sum(lens) * row_bytes
            self.bytes_written += out.numel() * out.element_size()
            if k == gate_from:
                gate_len = (od[1:] - od[:-1]).contiguous()
                self.keep.append(gate_len)
        self.gate = None
        if gate_len_dev is not None:   # frame counts already on the device (t2_melspec's n_frames): gate (B, gate_T, 1)
            assert gate_from is None and gate_len_dev.numel() == B
            gate_len = gate_len_dev.to(device=device, dtype=torch.int64).contiguous()
            self.keep.append(gate_len)
            self.gate = torch.empty((B, gate_T, 1), dtype=torch.float32, device=device)
            self.outs["gate"] = self.gate
            self.bytes_written += self.gate.numel() * 4
        if gate_from is not None:
            assert gate_len is not None, f"gate_from={gate_from!r} is not a field"
            self.gate = torch.empty((B, self.outs[gate_from].shape[1], 1), dtype=torch.float32, device=device)
            self.outs["gate"] = self.gate
            self.bytes_written += self.gate.numel() * 4
        self.gate_len = gate_len if self.gate is not None else None
        self.tables = (_table(C.c_void_p, src), _table(C.c_void_p, off), _table(C.c_longlong, rowb), _table(C.c_longlong, maxr),
                       _table(C.c_void_p, outp))
        self.n = len(names)

    def launch(self, stream=None):
        st = stream if stream is not None else torch.cuda.current_stream(self.device).cuda_stream
        t = [C.cast(x, C.c_void_p) for x in self.tables]
        _lib.check(self.lib.t2_collate(self.n, t[0], t[1], t[2], t[3], t[4], self.B, _lib.ptr(self.gate_len) if self.gate is not None else None,
                                       _lib.ptr(self.gate) if self.gate is not None else None,
                                       self.gate.shape[1] if self.gate is not None else 0, st), "t2_collate")
        return self.outs


def pad_fields(fields: dict, device, gate_from: str | None = None, stream=None):
    """fields: name -> list of B tensors (same dtype and trailing shape, ragged first dim).  Returns name -> (B, max, ...) device
    tensors; with ``gate_from`` = a field name also ``"gate"`` (B, max, 1) generated from that field's lengths."""
    return CollatePlan(fields, device, gate_from).launch(stream)


def collate(data, device="cuda:0"):
    """Drop-in for ``_collate`` (datasets/tts_dataloader.py:8-35) that assembles the batch on the GPU."""
    cols, lens, extra = defaultdict(list), defaultdict(list), defaultdict(list)
    for d, n, e in data:
        for k, v in d.items():
            cols[k].append(v)
        for k, v in n.items():
            lens[k].append(v)
        for k, v in e.items():
            extra[k].append(v)
    gen_gate = "gate" not in cols and "mel_spectrogram" in cols
    tts_data = pad_fields(dict(cols), device, gate_from="mel_spectrogram" if gen_gate else None)
    tts_data_len = {k: torch.stack(v).squeeze(1).to(device, non_blocking=True) for k, v in lens.items()}
    return tts_data, tts_data_len, dict(extra)


def trim_bounds(wav_pad, wav_len, top_db: float = 60.0, frame_length: int = 2048, hop_length: int = 512):
    """``librosa.effects.trim(wav, top_db=top_db, frame_length=frame_length)`` (datasets/tts_dataset.py:194-200) for a padded batch on
    the device: returns (start, end) int64 device tensors of the kept sample interval of every utterance (C ABI ``t2_trim``)."""
    lib = _lib.load()
    dev = wav_pad.device
    if dev.type != "cuda":
        raise RuntimeError("tacotron2_b200.collate: CUDA device required (no CPU fallback)")
    wav_pad = wav_pad.to(torch.float32).contiguous()
    wl = wav_len.to(device=dev, dtype=torch.int64).contiguous()
    B, stride = wav_pad.shape
    ws = torch.empty(lib.t2_trim_workspace_bytes(B, stride, hop_length), dtype=torch.uint8, device=dev)
    start, end = torch.empty(B, dtype=torch.int64, device=dev), torch.empty(B, dtype=torch.int64, device=dev)
    _lib.check(lib.t2_trim(_lib.ptr(wav_pad), _lib.ptr(wl), B, stride, stride, float(top_db), int(frame_length), int(hop_length), _lib.ptr(start),
                           _lib.ptr(end), _lib.ptr(ws), ws.numel(), torch.cuda.current_stream(dev).cuda_stream), "t2_trim")
    return start, end


class MelCache:
    """The reference's per-item mel cache (datasets/tts_dataset.py:184-188,206-207): ``cache_dir/<filename with '/' -> '_'>.pt`` holding
    the (T, n_mels) log-mel tensor written with ``torch.save``.  Files written here are read by the reference and vice versa."""

    def __init__(self, cache_dir: str):
        import os
        self.cache_dir = cache_dir
        os.makedirs(cache_dir, exist_ok=True)

    def path(self, filename: str) -> str:
        import os
        return os.path.join(self.cache_dir, f"{filename.replace('/', '_')}.pt")

    def has(self, filename: str) -> bool:
        import os
        return os.path.exists(self.path(filename))

    def load(self, filename: str, device=None):
        mel = torch.load(self.path(filename), map_location="cpu")
        return mel if device is None else mel.to(device, non_blocking=True)

    def save(self, filename: str, mel) -> None:
        torch.save(mel.detach().to("cpu", torch.float32).contiguous().clone(), self.path(filename))


def assemble_items(filenames, load_wav, chars, device="cuda:0", cache: "MelCache | None" = None, trim: bool = False, trim_top_db: float = 60.0,
                   trim_frame_length: int = 2048, silence: int = 0, melspectrogram=None):
    """``TTSDataset.__getitem__`` + ``_collate`` for a batch of files (datasets/tts_dataset.py:173-214, datasets/tts_dataloader.py:8-35)
    with the arithmetic on the GPU: items found in the mel cache are read from it, the others are loaded with ``load_wav(filename)``
    (1-D float tensor), trimmed / padded / turned into log-mels in one batch on the device and written back to the cache in the
    reference's format.  Returns ``(tts_data, tts_data_len)`` like ``assemble_from_wavs``."""
    device = torch.device(device)
    B = len(filenames)
    assert B == len(chars) and B > 0
    mels = [None] * B
    miss = [i for i, f in enumerate(filenames) if cache is None or not cache.has(f)]
    for i in range(B):
        if i not in miss:
            mels[i] = cache.load(filenames[i], device)
    if miss:
        data, lens = assemble_from_wavs([load_wav(filenames[i]) for i in miss], [chars[i] for i in miss], device, silence, melspectrogram, trim,
                                        trim_top_db, trim_frame_length)
        n = lens["mel_spectrogram_len"].tolist()
        for j, i in enumerate(miss):
            mels[i] = data["mel_spectrogram"][j, : n[j]]
            if cache is not None:
                cache.save(filenames[i], mels[i])
    out = pad_fields({"chars_idx": [c.to(torch.int64) for c in chars], "mel_spectrogram": [m.to(torch.float32) for m in mels]}, device,
                     gate_from="mel_spectrogram")
    nf = torch.tensor([int(m.shape[0]) for m in mels], dtype=torch.int32, device=device)
    return out, {"chars_idx_len": torch.tensor([int(c.numel()) for c in chars], dtype=torch.int64, device=device), "mel_spectrogram_len": nf,
                 "gate_len": nf.clone()}


def assemble_from_wavs(wavs, chars, device="cuda:0", silence: int = 0, melspectrogram=None, trim: bool = False, trim_top_db: float = 60.0,
                       trim_frame_length: int = 2048):
    """The GPU form of the per-item work of ``TTSDataset.__getitem__`` + ``_collate`` for a batch of items that are not in the mel
    cache (datasets/tts_dataset.py:194-214, datasets/tts_dataloader.py:8-35): waveforms (list of 1-D float tensors)
    are optionally trimmed (``trim=True``: ``librosa.effects.trim`` as :194-200, C ABI ``t2_trim``), padded with ``silence`` zeros
    (``F.pad(wav, (0, silence))``, :201), log-mels extracted for the whole batch by one
    ``t2_melspec`` launch straight into the padded (B, T, n_mels) layout, characters padded and gate targets written by one
    ``t2_collate`` launch.  Returns ``(tts_data, tts_data_len)`` with ``chars_idx``, ``mel_spectrogram``, ``gate`` and
    ``chars_idx_len`` (int64), ``mel_spectrogram_len`` / ``gate_len`` (int32, as the reference's IntTensor)."""
    from .melspec import HOP, TacotronMelSpectrogram

    device = torch.device(device)
    if device.type != "cuda":
        raise RuntimeError("tacotron2_b200.collate: CUDA device required (no CPU fallback)")
    mel_tr = melspectrogram if melspectrogram is not None else TacotronMelSpectrogram()
    B = len(wavs)
    assert B == len(chars) and B > 0
    if trim:   # librosa.effects.trim on the device (datasets/tts_dataset.py:194-200): bounds for the whole batch, one host read
        raw = pad_fields({"wav": [w.reshape(-1, 1).to(torch.float32) for w in wavs]}, device)["wav"].squeeze(-1)
        s_dev, e_dev = trim_bounds(raw, torch.tensor([int(w.numel()) for w in wavs], dtype=torch.int64), trim_top_db, trim_frame_length)
        se = torch.stack([s_dev, e_dev]).tolist()
        wavs = [raw[b, se[0][b]: se[1][b]] for b in range(B)]
    wl = torch.tensor([int(w.numel()) + silence for w in wavs], dtype=torch.int64)
    # ragged waveforms -> (B, S_max) zero padded on the device (the zero tail of each row is the silence and the batch padding at once)
    wav_pad = pad_fields({"wav": [w.reshape(-1, 1).to(torch.float32) for w in wavs]}, device)["wav"].squeeze(-1)
    s_max = int(wl.max())
    if wav_pad.shape[1] < s_max:
        wav_pad = torch.nn.functional.pad(wav_pad, (0, s_max - wav_pad.shape[1]))
    mel, n_frames = mel_tr.to(device).batch(wav_pad, wl)
    plan = CollatePlan({"chars_idx": [c.to(torch.int64) for c in chars]}, device, gate_len_dev=n_frames, gate_T=mel.shape[1])
    out = plan.launch()
    t_max = int(1 + s_max // HOP)
    assert mel.shape[1] == t_max
    data = {"chars_idx": out["chars_idx"], "mel_spectrogram": mel, "gate": out["gate"]}
    lens = {"chars_idx_len": torch.tensor([int(c.numel()) for c in chars], dtype=torch.int64, device=device),
            "mel_spectrogram_len": n_frames.to(torch.int32), "gate_len": n_frames.to(torch.int32)}
    return data, lens


def TTSDataLoader(dataset, batch_size=1, num_workers=0, pin_memory=False, pin_memory_device="", shuffle=None, prefetch_factor=None,
                  persistent_workers=False, drop_last=True, device="cuda:0"):
    """datasets/tts_dataloader.py:38-60 with the padding moved to the GPU: the worker processes only return the item triples
    (no pad_sequence on the host), the batch is assembled by ``collate`` in the consuming process."""
    from torch.utils.data import DataLoader

    loader = DataLoader(dataset, batch_size=batch_size, collate_fn=(lambda items: items) if batch_size > 1 else None,
                        num_workers=num_workers, pin_memory=pin_memory, pin_memory_device=pin_memory_device, shuffle=shuffle,
                        prefetch_factor=prefetch_factor if num_workers > 0 else None, persistent_workers=persistent_workers,
                        drop_last=drop_last)
    if batch_size <= 1:
        return loader

    class _GpuBatches:
        def __len__(self):
            return len(loader)

        def __iter__(self):
            for items in loader:
                yield collate(items, device)

    return _GpuBatches()
