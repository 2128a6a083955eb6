"""TEST INFRASTRUCTURE ONLY (checker for tests/, never imported by the product path).

CPU restatement of the reference's batch assembly:
  * ``_collate``            datasets/tts_dataloader.py:8-35  (pad_sequence over every data key, stack + squeeze of the lengths,
                                                            the "extra" dict passed through as lists)
  * gate targets            datasets/tts_dataset.py:211-214 (ones (T, 1), last frame 0)
Parity status: pinned -- tests/test_collate.py checks it against the reference's own ``_collate`` imported from /root/reference
when that tree is present, and against tests/golden/collate_golden.pt (generated from the reference by
tests/golden/make_collate_golden.py).
"""
from collections import defaultdict

import torch


def gate_target(n_frames: int) -> torch.Tensor:
    """datasets/tts_dataset.py:211-214."""
    g = torch.ones(n_frames, 1)
    g[-1] = 0.0
    return g


def pad_rows(items):
    """nn.utils.rnn.pad_sequence(batch_first=True) written out: (B, max_len, *trailing), zeros beyond each item's length."""
    mx = max(int(v.shape[0]) for v in items)
    out = torch.zeros((len(items), mx) + tuple(items[0].shape[1:]), dtype=items[0].dtype)
    for b, v in enumerate(items):
        out[b, : v.shape[0]] = v
    return out


def collate(data):
    """datasets/tts_dataloader.py:8-35 for a list of (data, lengths, extra) dict triples."""
    cols, lens, extra = defaultdict(list), defaultdict(list), defaultdict(list)
    for d, n, e in data:
        for k, v in d.items():
            cols[k].append(v)
        for k, v in n.items():
            lens[k].append(v)
        for k, v in e.items():
            extra[k].append(v)
    return ({k: pad_rows(v) for k, v in cols.items()}, {k: torch.stack(v).squeeze(1) for k, v in lens.items()}, dict(extra))
