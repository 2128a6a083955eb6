"""AR decode phase probe (profiling aid): C2 shapes, prints us/frame per kernel phase.  Env switches are read by the library."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import _lib, synth  # noqa: E402
from tacotron2_b200.model import Tacotron2  # noqa: E402

NAMES = ["att_tail", "qpart+bar", "attention", "dec_tail", "proj(side)", "pre1(side)", "pre2(side)", "xbar+kb0", "att_gemm", "dec_gemm", "a_qsum",
         "a_energy", "a_csync", "a_softmax", "a_ctx", "precompute(side)"]


def main():
    B, L, T = 64, 200, int(os.environ.get("PROBE_T", "400"))
    dev = torch.device("cuda", 0)
    a = synth.VANILLA
    model = Tacotron2(**a.kwargs()).to(dev).eval()
    model.load_state_dict(synth.synth_state_dict(a, seed=0, gate_shift=5.0))
    batch = synth.synth_batch(a, B, L, 8, seed=1234)
    c, n = batch["chars_idx"].to(dev), batch["chars_idx_len"].to(dev)
    with torch.no_grad():
        for _ in range(3):
            model(c, n, False, max_len_override=T)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        model._prof = []
        e0.record()
        out = model(c, n, False, max_len_override=T)[1]
        e1.record()
        e1.synchronize()
    dec = sum(x.elapsed_time(y) for x, y in model._prof)
    cyc = (ctypes.c_ulonglong * 16)()
    _lib.check(_lib.load().t2_debug_phase_cycles(model._ctx, cyc))
    ph = {k: round(cyc[i] / 1965 / T, 2) for i, k in enumerate(NAMES)}
    print("us/frame %.2f (frames %d)" % (dec * 1e3 / T, out.shape[1]), ph, flush=True)


if __name__ == "__main__":
    main()
