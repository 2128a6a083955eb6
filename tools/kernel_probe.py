"""Workload for the ncu metric pass over the tensor-core kernels off the recurrence: one Postnet.forward at the
C2 size (conv_tc_kernel x5: layer 2-4 are the 512 -> 512 convs) and one training step at C3 (gemm_nt_kernel: the bulk weight-gradient
GEMMs; bilstm_cluster_kernel / bilstm_bwd_cluster_kernel).  Run plain first, then under ncu with -k regex:'conv_tc|gemm_nt|bilstm'."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tacotron2_b200 import synth  # noqa: E402
from tacotron2_b200.model import Tacotron2  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    m = Tacotron2(**synth.VANILLA.kwargs()).to(dev).eval()
    m.load_state_dict(synth.synth_state_dict(synth.VANILLA, seed=0))
    x = torch.randn(64, 80, 1000, device=dev)
    with torch.no_grad():
        y = m.postnet(x)
    torch.cuda.synchronize()
    a = synth.HIFI
    t = Tacotron2(**a.kwargs()).to(dev).train()
    t.load_state_dict(synth.synth_state_dict(a, seed=0))
    b = {k: v.to(dev) for k, v in synth.synth_batch(a, 64, 150, 800, seed=1234, ragged=False).items()}
    mels, post, gates, _ = t(b["chars_idx"], b["chars_idx_len"], True, b["mel_spectrogram"], b["mel_spectrogram_len"], speaker_id=b["speaker_id"])
    loss = t.loss(mels, post, gates, b["mel_spectrogram"], b["gate"])
    loss.backward()
    torch.cuda.synchronize()
    t.check_status()
    print("probe ok", tuple(y.shape), float(loss))


if __name__ == "__main__":
    main()
