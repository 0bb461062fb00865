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


@pytest.mark.parametrize("rows,N,K", [(16, 32, 32), (1000, 128, 192), (4096, 256, 256), (300_000, 256, 192), (5, 64, 64),
                                      (300_001, 256, 256), (70_000, 192, 96)])
def test_wgrad_3xtf32_matches_float64(rows, N, K):
    """The 3xTF32 weight gradient (both streamed operands split in shared memory): FP32-level accuracy -- no worse
    than an FP32 cuBLAS product of the same operands, and ~1000x closer to float64 than the TF32 kernel's bar."""
    from gsb200.linear import linear_wgrad
    g = torch.Generator().manual_seed(rows % 89)
    go = torch.randn(rows, N, generator=g).cuda()
    x = torch.randn(rows, K, generator=g).cuda()
    dw = linear_wgrad(go, x, precise=True)
    torch.cuda.synchronize()
    want = go.double().t() @ x.double()
    scale = go.double().norm(dim=0)[:, None] * x.double().norm(dim=0)[None]
    err = ((dw.double() - want).abs() / scale).max().item()
    assert err < 4e-6, err
    acc = linear_wgrad(go, x, out=dw.clone(), accumulate=True, precise=True)
    assert ((acc.double() - 2 * want).abs() / scale).max().item() < 8e-6


@pytest.mark.parametrize("rows,M,H", [(1, 7, 256), (63, 7, 64), (5000, 7, 128), (300_000, 7, 256), (1000, 8, 32), (777, 3, 512)])
def test_thin_layer_backward_matches_float64(rows, M, H):
    from gsb200.linear import thin_backward
    g = torch.Generator().manual_seed(rows % 97)
    go = torch.randn(rows, M, generator=g).cuda()
    w = torch.randn(M, H, generator=g).cuda()
    x = torch.randn(rows, H, generator=g).cuda()
    dx, dw, db = thin_backward(go, w, x)
    torch.cuda.synchronize()
    want_dx = go.double() @ w.double()
    assert ((dx.double() - want_dx).abs().max() / want_dx.abs().max()).item() < 2e-6
    want_dw = go.double().t() @ x.double()
    scale = go.double().norm(dim=0)[:, None] * x.double().norm(dim=0)[None] + 1e-30
    assert ((dw.double() - want_dw).abs() / scale).max().item() < 4e-6
    want_db = go.double().sum(0)
    assert ((db.double() - want_db).abs().max() / (go.double().norm(dim=0).max() * rows ** 0.5 + 1e-30)).item() < 4e-6


@pytest.mark.parametrize("precise", [False, True])
@pytest.mark.parametrize("rows,K,N", [(1000, 192, 128), (300_001, 256, 256), (5, 64, 64)])
def test_linear_with_residual_added_in_the_epilogue(rows, K, N, precise):
    from gsb200.linear import linear
    x, w, _, want, scale = _case(rows, K, N, seed=rows % 97, bias=False)
    res = torch.randn(rows, N, generator=torch.Generator().manual_seed(1)).cuda()
    y = linear(x, w, None, precise=precise, residual=res)
    torch.cuda.synchronize()
    err = ((y.double() - (want + res.double())).abs() / (scale + res.double().abs())).max().item()
    assert err < (2e-6 if precise else 2e-3), err
    inplace = res.clone()
    linear(x, w, None, precise=precise, residual=inplace, out=inplace)     # Y may alias the residual
    assert torch.equal(inplace, y)
