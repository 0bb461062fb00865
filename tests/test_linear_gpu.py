"""tcgen05 dense layer against a float64 torch matmul.  Tolerances, per output element and
relative to |x_row| |w_col| (the scale of the products that are rounded): TF32 2^-10 per operand -> 2e-3; 3xTF32 -> 2e-6."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(rows, K, N, seed, bias=True):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(rows, K, generator=g).cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).cuda()
    b = torch.randn(N, generator=g).cuda() if bias else None
    want = x.double() @ w.double().t() + (b.double() if bias else 0)
    scale = x.double().norm(dim=1, keepdim=True) * w.double().norm(dim=1)[None]
    return x, w, b, want, scale


@pytest.mark.parametrize("rows,K,N", [(128, 32, 16), (1000, 192, 128), (4096, 256, 256), (777, 128, 7), (300_000, 192, 256),
                                      (5, 64, 64), (300_001, 256, 128)])
@pytest.mark.parametrize("precise", [False, True])
def test_linear_matches_float64(rows, K, N, precise):
    from gsb200.linear import linear
    x, w, b, want, scale = _case(rows, K, N, seed=rows % 97)
    y = linear(x, w, b, precise=precise)
    torch.cuda.synchronize()
    err = ((y.double() - want).abs() / scale).max().item()
    assert err < (2e-6 if precise else 2e-3), err
    assert y.shape == (rows, N) and y.dtype == torch.float32


def test_linear_no_bias_and_strided_input():
    from gsb200.linear import linear
    x, w, _, want, scale = _case(2048, 64, 32, seed=3, bias=False)
    wide = torch.zeros(2048, 96, device="cuda")
    wide[:, :64] = x
    y = linear(wide[:, :64], w, None)     # row pitch 96 floats: no copy needed
    assert ((y.double() - want).abs() / scale).max().item() < 2e-3


@pytest.mark.parametrize("rows,N,K", [(32, 128, 32), (1000, 128, 192), (4096, 256, 256), (300_000, 256, 192), (5, 64, 64),
                                      (300_001, 128, 128), (70_000, 192, 96)])
def test_wgrad_matches_float64(rows, N, K):
    from gsb200.linear import linear_wgrad
    g = torch.Generator().manual_seed(rows % 89)
    go = torch.randn(rows, N, generator=g).cuda()
    x = torch.randn(rows, K, generator=g).cuda()
    dw = linear_wgrad(go, x)
    torch.cuda.synchronize()
    want = go.double().t() @ x.double()
    scale = go.double().norm(dim=0)[:, None] * x.double().norm(dim=0)[None]
    err = ((dw.double() - want).abs() / scale).max().item()
    assert err < 2e-3, err
    acc = linear_wgrad(go, x, out=dw.clone(), accumulate=True)
    assert ((acc.double() - 2 * want).abs() / scale).max().item() < 4e-3
