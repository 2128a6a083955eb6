"""Phase profile of the teacher-forced training forward (decoder_ar_kernel<true>) at C3: us per frame per kernel phase (CTA 0)."""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import _lib, synth  # noqa: E402
from tacotron2_b200.model import Tacotron2  # noqa: E402

NAMES = ["att_tail", "qpart+Qbar", "Cbar+toeplitz", "dec_tail+Rbar", "proj", "prenet1", "prenet2", "Xbar+stop", "wait_acc_att", "wait_acc_dec",
         "a_qsum+locwait", "a_energy", "a_exchange", "a_softmax", "a_ctx", "proj_ctx"]
dev = torch.device("cuda:0")
a = synth.HIFI
T = int(os.environ.get("PROBE_T", "800"))
model = Tacotron2(**a.kwargs()).to(dev).train()
model.load_state_dict(synth.synth_state_dict(a, seed=0))
b = {k: v.to(dev) for k, v in synth.synth_batch(a, 64, 150, T, seed=1234, ragged=False).items()}
lib = _lib.load()
for it in range(3):
    out = model(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"], speaker_id=b["speaker_id"])
    torch.cuda.synchronize()
    if it == 1:
        _lib.check(lib.t2_debug_set_timing(model._ctx, 1))
v = ctypes.c_float(0.0)
_lib.check(lib.t2_debug_kernel_ms(model._ctx, 0, ctypes.byref(v)))
cyc = (ctypes.c_ulonglong * 16)()
_lib.check(lib.t2_debug_phase_cycles(model._ctx, cyc))
print(f"TF forward kernel {v.value:.2f} ms = {1e3 * v.value / T:.2f} us/frame", {k: round(cyc[i] / 1965 / T, 2) for i, k in enumerate(NAMES)})
