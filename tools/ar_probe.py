"""A/B probe of the two autoregressive decoder schedules (profiling / bring-up aid): the pipelined kernel (decoder_ar.cu, default) against
decoder_persistent_kernel (T2B200_AR_KERNEL=old) on the same inputs: max-abs output differences, us per frame, per-phase cycles.
usage: ar_probe.py [B L T] ..."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import _lib, synth  # noqa: E402
from tacotron2_b200.model import Tacotron2  # noqa: E402

NEW_NAMES = ["att_tail", "qpart+Qbar", "Cbar+toeplitz", "dec_tail+Rbar", "proj", "prenet1", "prenet2", "Xbar+stop", "wait_acc_att", "wait_acc_dec",
             "a_qsum+locwait", "a_energy", "a_exchange", "a_softmax", "a_ctx", "proj_ctx"]


def run(model, c, n, T, masks, kind, reps=2):
    if kind == "old":
        os.environ["T2B200_AR_KERNEL"] = "old"
    else:
        os.environ.pop("T2B200_AR_KERNEL", None)
    lib = _lib.load()
    with torch.no_grad():
        out = model(c, n, False, max_len_override=T, dropout_masks=masks)
        _lib.check(lib.t2_debug_set_timing(model._ctx, 1))
        ms = []
        for _ in range(reps):
            out = model(c, n, False, max_len_override=T, dropout_masks=masks)
            torch.cuda.synchronize()
            v = ctypes.c_float(0.0)
            _lib.check(lib.t2_debug_kernel_ms(model._ctx, 0, ctypes.byref(v)))
            ms.append(v.value)
    cyc = (ctypes.c_ulonglong * 16)()
    _lib.check(lib.t2_debug_phase_cycles(model._ctx, cyc))
    model.check_status()
    return out, min(ms), [cyc[i] for i in range(16)]


def main():
    shapes = []
    a = sys.argv[1:]
    while len(a) >= 3:
        shapes.append((int(a[0]), int(a[1]), int(a[2])))
        a = a[3:]
    if not shapes:
        shapes = [(4, 37, 12), (64, 200, 200)]
    dev = torch.device("cuda", 0)
    args = synth.VANILLA
    for B, L, T in shapes:
        model = Tacotron2(**args.kwargs()).to(dev).eval()
        model.load_state_dict(synth.synth_state_dict(args, seed=0, gate_shift=5.0))
        batch = synth.synth_batch(args, B, L, 8, seed=1234)
        masks = synth.synth_masks(args, B, L, T, seed=7)
        c, n = batch["chars_idx"].to(dev), batch["chars_idx_len"].to(dev)
        new, ms_new, cyc = run(model, c, n, T, masks, "new")
        old, ms_old, _ = run(model, c, n, T, masks, "old")
        diffs = {k: float((x - y).abs().max()) for k, x, y in zip(("mel", "post", "gate", "align"), new, old)}
        print(f"B={B} L={L} T={T}: new {1e3 * ms_new / T:.2f} us/frame, old {1e3 * ms_old / T:.2f} us/frame, frames {new[0].shape[1]}/{old[0].shape[1]}, diff {diffs}", flush=True)
        print("   new phases us/frame:", {k: round(cyc[i] / 1965 / T, 2) for i, k in enumerate(NEW_NAMES)}, flush=True)
        del model


if __name__ == "__main__":
    main()
