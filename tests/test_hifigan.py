"""HiFi-GAN generator through the C ABI against the UNMODIFIED reference generator (model/hifi_gan.py:154-216):
committed goldens (tests/golden/make_golden_hifigan.py) and, when oracle/_ref is staged, the reference module itself on larger
shapes.  Bar: max-abs waveform error <= 1e-3; the waveform is tanh-bounded in [-1, 1]."""
import contextlib
import io
import os

import numpy as np
import pytest
import torch

from tacotron2_b200 import hifigan as H

G = os.path.join(os.path.dirname(__file__), "golden")
WAV_TOL = 1e-3
V3LIKE = dict(H.UNIVERSAL_V1, resblock="2", upsample_initial_channel=256, upsample_rates=[8, 8, 4], upsample_kernel_sizes=[16, 16, 8],
              resblock_kernel_sizes=[3, 5, 7], resblock_dilation_sizes=[[1, 2], [2, 6], [3, 12]])


class AttrDict(dict):
    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.__dict__ = self


def golden_mel(B, T, seed):
    g = np.random.Generator(np.random.PCG64(seed))
    return torch.from_numpy(np.clip(g.standard_normal((B, 80, T)) * 2.0 - 5.0, np.log(1e-5), 2.0).astype(np.float32))


def test_state_dict_keys_and_weight_norm_folding_cpu():
    """Same parameter names / shapes as the reference generator after remove_weight_norm(); weight_g / weight_v checkpoints fold."""
    g = H.Generator(AttrDict(H.UNIVERSAL_V1))
    sd = g.state_dict()
    assert sd["conv_pre.weight"].shape == (512, 80, 7) and sd["ups.0.weight"].shape == (512, 256, 16) and sd["conv_post.weight"].shape == (1, 32, 7)
    assert sd["resblocks.11.convs2.2.weight"].shape == (32, 32, 11) and len(sd) == 2 * (2 + 4 + 12 * 6)
    assert abs(H.generator_flops_per_frame(H.UNIVERSAL_V1) / 1e6 - 600) < 40
    from oracle import ref_model as R
    if R.available():
        with contextlib.redirect_stdout(io.StringIO()):
            ref = R.reference_generator(AttrDict(H.UNIVERSAL_V1))
            wn = {k: v.clone() for k, v in ref.state_dict().items()}     # weight-normalised form (weight_g / weight_v)
            ref.remove_weight_norm()
        assert sorted(ref.state_dict().keys()) == sorted(sd.keys())   # (remove_weight_norm re-registers weight after bias: same set)
        assert all(ref.state_dict()[k].shape == sd[k].shape for k in sd)
        g.load_state_dict(wn)
        for k, v in ref.state_dict().items():
            assert float((g.state_dict()[k] - v).abs().max()) <= 1e-6 * max(1.0, float(v.abs().max())), k


@pytest.mark.gpu
@pytest.mark.parametrize("name,cfg", [("hifigan_v1", H.UNIVERSAL_V1), ("hifigan_v3like", V3LIKE)])
def test_generator_golden(name, cfg):
    g = dict(np.load(os.path.join(G, name + ".npz")))
    B, T, seed = (int(x) for x in g["meta"])
    dev = torch.device("cuda:0")
    m = H.Generator(AttrDict(cfg)).to(dev).eval()
    m.load_state_dict(H.synth_generator_state(cfg, seed=0))
    m.remove_weight_norm()
    with torch.no_grad():
        wav = m(golden_mel(B, T, seed).to(dev))
    assert tuple(wav.shape) == g["wav"].shape
    err = float(np.abs(wav.cpu().numpy() - g["wav"]).max())
    assert err <= WAV_TOL, err
    assert float(np.abs(g["wav"]).std()) > 0.05, "golden is degenerate"


@pytest.mark.gpu
@pytest.mark.parametrize("B,T", [(1, 1), (3, 7), (2, 200), (8, 800)])
def test_generator_vs_staged_reference(B, T):
    """Against the reference module itself (oracle/_ref) on the host: single frame, ragged tile edges, and the 8 x 800-frame shape of
    VERDICT r1 item 7 (compared on the first two utterances there: the CPU reference needs ~0.5 s per 100 frames)."""
    from oracle import ref_model as R
    if not R.available():
        pytest.skip("oracle/_ref not staged")
    torch.set_num_threads(os.cpu_count() or 1)
    dev = torch.device("cuda:0")
    sd = H.synth_generator_state(H.UNIVERSAL_V1, seed=1)
    m = H.Generator(AttrDict(H.UNIVERSAL_V1)).to(dev).eval()
    m.load_state_dict(sd)
    x = golden_mel(B, T, 11)
    with torch.no_grad():
        wav = m(x.to(dev))
    assert tuple(wav.shape) == (B, 1, 256 * T) and bool(torch.isfinite(wav).all())
    with contextlib.redirect_stdout(io.StringIO()):
        ref = R.reference_generator(AttrDict(H.UNIVERSAL_V1))
        ref.remove_weight_norm()
    ref.load_state_dict(sd)
    ref.eval()
    nb = min(B, 2)
    with torch.no_grad():
        want = ref(x[:nb])
    err = float((wav[:nb].cpu() - want).abs().max())
    assert err <= WAV_TOL, err
