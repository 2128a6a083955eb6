"""Per-stage CUDA-event timing of the training step's forward and backward (which autograd Function the time goes to), several steps in
a row, to expose run-to-run variation.  usage: python tools/train_probe.py [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import autograd as AG  # noqa: E402
from tacotron2_b200 import synth  # noqa: E402
from tacotron2_b200.model import Tacotron2  # noqa: E402
from tacotron2_b200.optim import ClipAdam  # noqa: E402

dev = torch.device("cuda:0")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
spans = []


def wrap(cls):
    f0, b0 = cls.forward, cls.backward

    def fwd(ctx, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = f0(ctx, *a); e1.record()  # noqa: E702
        spans.append((cls.__name__ + ".fwd", e0, e1))
        return out

    def bwd(ctx, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); out = b0(ctx, *a); e1.record()  # noqa: E702
        spans.append((cls.__name__ + ".bwd", e0, e1))
        return out

    cls.forward, cls.backward = staticmethod(fwd), staticmethod(bwd)


for c in (AG.PostnetFn, AG.EncoderFn, AG.MemoryFn, AG.DecodeTfFn, AG.FinalizeFn, AG.LossFn):
    wrap(c)
a = synth.HIFI
model = Tacotron2(**a.kwargs()).to(dev).train()
model.load_state_dict(synth.synth_state_dict(a, seed=0))
if os.environ.get("PROBE_PREALLOC"):   # something large allocated and freed first, as bench.py's AR half does
    x = torch.empty(int(os.environ["PROBE_PREALLOC"]) << 20, dtype=torch.uint8, device=dev); x.zero_(); del x  # noqa: E702
    torch.cuda.empty_cache()
b = {k: v.to(dev) for k, v in synth.synth_batch(a, 64, 150, 800, seed=1234, ragged=False).items()}
opt = ClipAdam(model.parameters(), lr=1e-3, weight_decay=1e-6)
for s in range(steps):
    spans.clear()
    opt.zero_grad()
    t0, t1, t2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    t0.record()
    mels, post, gates, _ = model(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"], speaker_id=b["speaker_id"])
    loss = model.loss(mels, post, gates, b["mel_spectrogram"], b["gate"])
    t1.record()
    loss.backward()
    t2.record()
    opt.step()
    torch.cuda.synchronize()
    print(f"step {s}: fwd {t0.elapsed_time(t1):7.2f}  bwd {t1.elapsed_time(t2):7.2f}  | " +
          "  ".join(f"{n} {e0.elapsed_time(e1):.2f}" for n, e0, e1 in spans), flush=True)
print("mem reserved MB", torch.cuda.memory_reserved() >> 20, "allocated", torch.cuda.memory_allocated() >> 20)
