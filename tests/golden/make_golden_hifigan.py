"""Golden waveform of the UNMODIFIED reference HiFi-GAN generator (run in the build container only).

    python tests/golden/make_golden_hifigan.py        # needs /root/reference

Builds model/hifi_gan.py:Generator with the UNIVERSAL_V1 configuration run/say.py:76-84 loads, removes weight norm as
run/say.py:88 does, loads the seeded weights of tacotron2_b200.hifigan.synth_generator_state and stores its output for a small
seeded mel batch (inputs are regenerated from the seed by the test)."""
import contextlib
import io
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.abspath(os.path.join(HERE, "..", "..")))
sys.path.insert(0, "/root/reference")

from tacotron2_b200 import hifigan as H  # noqa: E402

with contextlib.redirect_stdout(io.StringIO()):
    from model.hifi_gan import Generator  # noqa: E402  (the reference)


class AttrDict(dict):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.__dict__ = self


def golden_mel(B, T, seed):
    g = np.random.Generator(np.random.PCG64(seed))
    return torch.from_numpy(np.clip(g.standard_normal((B, 80, T)) * 2.0 - 5.0, np.log(1e-5), 2.0).astype(np.float32))


if __name__ == "__main__":
    torch.set_num_threads(8)
    for name, cfg, B, T in (("hifigan_v1", H.UNIVERSAL_V1, 2, 24),
                            ("hifigan_v3like", dict(H.UNIVERSAL_V1, resblock="2", upsample_initial_channel=256, upsample_rates=[8, 8, 4], upsample_kernel_sizes=[16, 16, 8],
                                                    resblock_kernel_sizes=[3, 5, 7], resblock_dilation_sizes=[[1, 2], [2, 6], [3, 12]]), 1, 16)):
        with contextlib.redirect_stdout(io.StringIO()):
            ref = Generator(AttrDict(cfg))
            ref.remove_weight_norm()
        ref.load_state_dict(H.synth_generator_state(cfg, seed=0), strict=True)
        ref.eval()
        with torch.no_grad():
            wav = ref(golden_mel(B, T, 5))
        print(name, tuple(wav.shape), float(wav.abs().max()), float(wav.std()))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), wav=wav.numpy(), meta=np.array([B, T, 5], np.int64))
