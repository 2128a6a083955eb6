"""Fused clip + Adam (t2_clip_adam_step) against torch.nn.utils.clip_grad_norm_ + torch.optim.Adam -- the optimizer shell the
reference configures in model/tts_model.py:78-91 and run/train.py:239-240."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _make(shapes, seed, dev):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter((torch.randn(s, generator=g) * 0.3).to(dev)) for s in shapes]


@pytest.mark.parametrize("wd,max_norm,gscale", [(1e-6, 1.0, 1.0), (0.0, 1.0, 1e-3), (1e-2, 0.5, 30.0)])
def test_clip_adam_matches_torch(wd, max_norm, gscale):
    from tacotron2_b200.optim import ClipAdam

    dev = torch.device("cuda:0")
    shapes = [(1,), (7, 3), (2048,), (2049,), (513, 129), (80, 1536), (31,)]   # chunk-boundary and ragged sizes
    pa, pb = _make(shapes, 0, dev), _make(shapes, 0, dev)
    ref = torch.optim.Adam(pa, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    opt = ClipAdam(pb, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd, max_norm=max_norm)
    g = torch.Generator().manual_seed(1)
    for step in range(5):
        grads = [(torch.randn(s, generator=g) * gscale).to(dev) for s in shapes]
        for a, b, gr in zip(pa, pb, grads):
            a.grad, b.grad = gr.clone(), gr.clone()
        n_ref = torch.nn.utils.clip_grad_norm_(pa, max_norm)
        ref.step()
        v0 = pb[0]._version
        n = opt.step()
        assert pb[0]._version > v0   # the model's operand re-pack watches the version counters
        assert abs(float(n) - float(n_ref)) <= 1e-5 * float(n_ref)
        if step == 2:
            for grp in ref.param_groups:
                grp["lr"] = 5e-4
            opt.set_lr(5e-4)
    for a, b in zip(pa, pb):
        # same formula in fp32; torch's foreach path orders a few roundings differently
        assert float((a - b).abs().max()) <= 2e-6 + 1e-5 * float(a.abs().max())
    st = ref.state[pa[4]]
    assert torch.allclose(st["exp_avg"], opt.exp_avg[4], rtol=1e-5, atol=1e-8)
    assert torch.allclose(st["exp_avg_sq"], opt.exp_avg_sq[4], rtol=1e-5, atol=1e-10)


def test_clip_adam_rejects_cpu():
    from tacotron2_b200.optim import ClipAdam

    with pytest.raises(RuntimeError):
        ClipAdam([torch.nn.Parameter(torch.zeros(4))])
