"""tcgen05 split-K GEMM (the decoder's gate-contraction engine) vs a float64 matmul."""
import pytest
import torch

from tacotron2_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,K,npass,tol", [(128, 256, 3, 2e-5), (256, 1792, 3, 2e-5), (128, 2560, 1, 8e-3), (4096, 2560, 3, 2e-5)])
def test_tcgemm(rows, K, npass, tol):
    dev = torch.device("cuda:0")
    lib = _lib.load()
    g = torch.Generator().manual_seed(rows + K)
    W = (torch.rand(rows, K, generator=g) * 2 - 1) / 32
    x = torch.randn(64, K, generator=g)
    Wd, xd = W.to(dev), x.to(dev)
    out = torch.full((rows, 64), float("nan"), device=dev)
    ws = torch.empty(4 * (rows + 64) * K + (1 << 16), dtype=torch.uint8, device=dev)
    _lib.check(lib.t2_debug_tcgemm(_lib.ptr(Wd), _lib.ptr(xd), _lib.ptr(out), rows, K, npass, _lib.ptr(ws), ws.numel(),
                                   torch.cuda.current_stream().cuda_stream), "t2_debug_tcgemm")
    want = (W.double() @ x.double().T).float()
    err = float((out.cpu() - want).abs().max())
    scale = float(want.abs().max())
    assert err <= tol * max(1.0, scale), (err, scale)
