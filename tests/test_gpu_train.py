"""GPU parity of the training path: every module's train-mode forward and backward, through the C ABI,
against torch autograd over the CPU oracle on the same seeded inputs and injected dropout masks.

Gradient tolerance: max-abs error <= 2e-3 * max|g| per tensor (the reference itself trains under TF32 + fp16 autocast,
run/train.py:170,239; the kernels here use three-pass bf16 (hi,lo) tensor-core arithmetic, ~16 mantissa bits)."""
import numpy as np
import pytest
import torch

from oracle import tacotron2_oracle as O
from tacotron2_b200 import synth
from tacotron2_b200.model import Tacotron2

pytestmark = pytest.mark.gpu
GTOL = 2e-3


def _dev():
    assert torch.cuda.is_available()
    return torch.device("cuda:0")


def _model(args, sd, train=True):
    m = Tacotron2(**args.kwargs()).to(_dev())
    m.load_state_dict(sd)
    return m.train(train)


def _sd_grad(sd):
    return {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running_" not in k else v.clone()) for k, v in sd.items()}


def _check_grads(model, sd, prefix, tol=GTOL):
    bad = []
    for k, p in model.named_parameters():
        if not k.startswith(prefix):
            continue
        want = sd[k].grad if sd[k].grad is not None else torch.zeros_like(sd[k])
        assert p.grad is not None, f"{k}: no gradient"
        got = p.grad.detach().cpu()
        scale = max(float(want.abs().max()), 1e-8)
        err = float((got - want).abs().max())
        if not np.isfinite(err) or err > tol * scale + 1e-9:
            bad.append((k, err, scale))
    assert not bad, bad


def test_loss_kernel():
    a = synth.VANILLA
    m = _model(a, synth.synth_state_dict(a))
    g = torch.Generator().manual_seed(3)
    B, T = 5, 37
    mels, post = torch.randn(B, T, 80, generator=g), torch.randn(B, T, 80, generator=g)
    gates = torch.randn(B, T, 1, generator=g) * 3
    gates[0, 30:] = -1000.0
    tgt = torch.randn(B, T, 80, generator=g)
    gt = (torch.rand(B, T, 1, generator=g) > 0.3).float()
    ins = [t.clone().requires_grad_(True) for t in (mels, post, gates)]
    want = O.loss_fn(ins[0], ins[1], ins[2], tgt, gt)
    want.backward()
    cu = [t.clone().to(_dev()).requires_grad_(True) for t in (mels, post, gates)]
    got = m.loss(cu[0], cu[1], cu[2], tgt.to(_dev()), gt.to(_dev()))
    got.backward()
    assert abs(float(got) - float(want)) <= 1e-5 * abs(float(want))
    for a_, b_ in zip(cu, ins):
        assert float((a_.grad.cpu() - b_.grad).abs().max()) <= 1e-7 + 1e-5 * float(b_.grad.abs().max())


@pytest.mark.parametrize("B,T", [(3, 17), (4, 200)])
def test_postnet_train_fwd_bwd(B, T):
    a = synth.VANILLA
    sd = synth.synth_state_dict(a)
    m = _model(a, sd)
    keep = synth.synth_masks(a, B, 5, T, seed=11, train=True)["postnet"]
    g = torch.Generator().manual_seed(5)
    x = torch.randn(B, 80, T, generator=g)
    dy = torch.randn(B, 80, T, generator=g)
    sdg = _sd_grad(sd)
    xo = x.clone().requires_grad_(True)
    want = O.postnet(sdg, xo, train=True, keep=keep, p=a.dropout)
    (want * dy).sum().backward()
    xc = x.clone().to(_dev()).requires_grad_(True)
    got = m.postnet(xc, dropout_keep=keep)
    (got * dy.to(_dev())).sum().backward()
    assert float((got.detach().cpu() - want.detach()).abs().max()) <= 2e-4
    scale = float(xo.grad.abs().max())
    assert float((xc.grad.cpu() - xo.grad).abs().max()) <= GTOL * scale
    _check_grads(m, sdg, "postnet.")
    # running statistics (momentum 0.1, unbiased variance), first layer
    z = torch.nn.functional.conv1d(x, sd["postnet.postnet.0.weight"], None, padding=2)
    rm = 0.9 * sd["postnet.postnet.1.running_mean"] + 0.1 * z.mean(dim=(0, 2))
    rv = 0.9 * sd["postnet.postnet.1.running_var"] + 0.1 * z.var(dim=(0, 2), unbiased=True)
    st = m.state_dict()
    assert float((st["postnet.postnet.1.running_mean"].cpu() - rm).abs().max()) <= 1e-5
    assert float((st["postnet.postnet.1.running_var"].cpu() - rv).abs().max()) <= 1e-5
