"""Dense layers on the tcgen05 GEMM of csrc/gemm.cu (the DeformationNetwork's nn.Linear
layers, train.py:57-110).  linear(x, weight, bias) == torch.nn.functional.linear(x, weight, bias) with TF32
multiplies (precise=False) or the 3xTF32 operand split (precise=True, FP32-level accuracy).  Forward only; CUDA only."""
import torch

from . import _lib as L

_scratch = {}


def linear(x, weight, bias=None, precise=False, out=None, col_stats=None, residual=None):
    """col_stats: optional zeroed float64 [2, N] tensor receiving the column sums / sums of squares of the output.
    residual: optional [rows, N] tensor added to the output in the GEMM epilogue (no bias, no col_stats then)."""
    L.require_cuda(x, "x")
    x = x if (x.dtype == torch.float32 and x.stride(-1) == 1 and x.dim() == 2) else x.reshape(-1, x.shape[-1]).float().contiguous()
    w = weight.detach().float().contiguous()
    b = None if bias is None else bias.detach().float().contiguous()
    rows, K = x.shape
    N = w.shape[0]
    if w.shape[1] != K:
        raise RuntimeError(f"linear: x has {K} columns, weight {tuple(w.shape)}")
    if x.stride(0) % 4 or x.data_ptr() % 16:
        x = x.contiguous()
        if K % 4:
            raise RuntimeError("linear: K must be a multiple of 4 (TMA row pitch)")
    y = out if out is not None else torch.empty((rows, N), dtype=torch.float32, device=x.device)
    if residual is not None:
        if bias is not None or col_stats is not None:
            raise RuntimeError("linear: residual excludes bias and col_stats")
        r = residual if (residual.dtype == torch.float32 and residual.stride(-1) == 1) else residual.float().contiguous()
        nbytes = L.lib().gsb_linear_tf32x3_scratch_bytes(K, N)
        key = (x.device, nbytes)
        if precise and key not in _scratch:
            _scratch[key] = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        with L.device_guard(x.device):
            rc = L.lib().gsb_linear_residual(rows, K, N, L.ptr(x), x.stride(0), L.ptr(w), K, L.ptr(r), r.stride(0), L.ptr(y),
                                             y.stride(0), 1 if precise else 0, L.ptr(_scratch[key]) if precise else None,
                                             L.stream_ptr())
        L.check(rc, "gsb_linear_residual")
        return y
    with L.device_guard(x.device):
        if precise:
            nbytes = L.lib().gsb_linear_tf32x3_scratch_bytes(K, N)
            key = (x.device, nbytes)
            if key not in _scratch:
                _scratch[key] = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
            rc = L.lib().gsb_linear_tf32x3(rows, K, N, L.ptr(x), x.stride(0), L.ptr(w), K, L.ptr(b), L.ptr(y), y.stride(0),
                                           L.ptr(col_stats), L.ptr(_scratch[key]), L.stream_ptr())
        else:
            rc = L.lib().gsb_linear_tf32(rows, K, N, L.ptr(x), x.stride(0), L.ptr(w), K, L.ptr(b), L.ptr(y), y.stride(0),
                                         L.ptr(col_stats), L.stream_ptr())
    L.check(rc, "gsb_linear_tf32")
    return y


def linear_wgrad(grad_out, x, out=None, accumulate=False, precise=False):
    """dW [N, K] = grad_out[rows, N]^T @ x[rows, K] on the tcgen05 weight-gradient kernel: TF32 multiplies, or the
    3xTF32 operand split (precise=True: FP32-level accuracy, what the reference's FP32 autograd computes)."""
    L.require_cuda(x, "x")
    g = grad_out if (grad_out.dtype == torch.float32 and grad_out.stride(-1) == 1) else grad_out.float().contiguous()
    x = x if (x.dtype == torch.float32 and x.stride(-1) == 1) else x.float().contiguous()
    rows, n = g.shape
    k = x.shape[1]
    if x.shape[0] != rows:
        raise RuntimeError("linear_wgrad: row counts differ")
    if n % 32 or k % 32 or n > 256 or k > 256:   # outside the tensor-core tiling: plain product
        dw = g.t() @ x
        if out is not None:
            out.copy_(out + dw if accumulate else dw)
            return out
        return dw
    dw = out if out is not None else torch.empty((n, k), dtype=torch.float32, device=x.device)
    nbytes = L.lib().gsb_linear_wgrad_scratch_bytes(n, k)
    key = (x.device, "wgrad", nbytes)
    if key not in _scratch:
        _scratch[key] = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    with L.device_guard(x.device):
        fn = L.lib().gsb_linear_wgrad_tf32x3 if precise else L.lib().gsb_linear_wgrad_tf32
        rc = fn(rows, n, k, L.ptr(g), g.stride(0), L.ptr(x), x.stride(0), L.ptr(dw), 1 if accumulate else 0,
                L.ptr(_scratch[key]), L.stream_ptr())
    L.check(rc, "gsb_linear_wgrad_tf32")
    return dw


def thin_backward(grad_out, weight, x, parts="both"):
    """The backward of a dense layer with at most 8 output columns (the network's 7-wide output layer,
    train.py:80-110) in FP32: -> (dX [rows, H] = grad_out @ weight, dW [M, H] = grad_out^T @ x, db [M])."""
    L.require_cuda(x, "x")
    g = grad_out if (grad_out.dtype == torch.float32 and grad_out.stride(-1) == 1) else grad_out.float().contiguous()
    x = x if (x.dtype == torch.float32 and x.stride(-1) == 1) else x.float().contiguous()
    w = weight.detach().float().contiguous()
    rows, m = g.shape
    h = w.shape[1]
    if w.shape[0] != m or x.shape != (rows, h):
        raise RuntimeError("thin_backward: shapes differ")
    dx = torch.empty((rows, h), dtype=torch.float32, device=x.device)
    dw = torch.empty((m, h), dtype=torch.float32, device=x.device)
    db = torch.empty((m,), dtype=torch.float32, device=x.device)
    nbytes = L.lib().gsb_linear_thin_wgrad_scratch_bytes(m, h)
    key = (x.device, "thin", nbytes)
    if key not in _scratch:
        _scratch[key] = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    with L.device_guard(x.device):
        if parts != "dw":   # (`parts` exists for the bench: one product alone)
            L.check(L.lib().gsb_linear_thin_dx(rows, m, h, L.ptr(g), g.stride(0), L.ptr(w), h, L.ptr(dx), dx.stride(0),
                                               L.stream_ptr()), "gsb_linear_thin_dx")
        if parts != "dx":
            L.check(L.lib().gsb_linear_thin_wgrad(rows, m, h, L.ptr(g), g.stride(0), L.ptr(x), x.stride(0), L.ptr(dw),
                                                  L.ptr(db), L.ptr(_scratch[key]), L.stream_ptr()), "gsb_linear_thin_wgrad")
    return dx, dw, db
