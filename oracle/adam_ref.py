"""TEST INFRASTRUCTURE (oracle): numpy restatement of the two elementwise steps either side of the rasterizer.

  activate / activate_backward   create_render_arguments, shared.py:29-42 (torch.nn.functional.normalize,
                                 torch.sigmoid, torch.exp and their autograd formulas)
  adam_step                      torch.optim.Adam as create_optimizer configures it (densify.py:68-86), the
                                 single-tensor arithmetic of torch/optim/adam.py (third-party: PyTorch, reference pin
                                 1.12.1, here 2.11.0): lerp of exp_avg, mul+addcmul of exp_avg_sq, bias corrections
                                 in Python doubles, addcdiv.
Pinned in tests/test_optim.py against torch itself (autograd of the three torch ops; torch.optim.Adam) on the
CPU.  Only tests may import this module.
"""
import numpy as np


def activate(quats, logits, log_scales):
    q = quats.astype(np.float32)
    n = np.sqrt((q * q).sum(-1, keepdims=True, dtype=np.float32))
    rot = q / np.maximum(n, np.float32(1e-12))
    opa = (np.float32(1) / (np.float32(1) + np.exp(-logits.astype(np.float32)))).astype(np.float32)
    return rot.astype(np.float32), opa, np.exp(log_scales.astype(np.float32))


def activate_backward(quats, d_rot, d_opa, d_sca, opa, sca):
    q = quats.astype(np.float32)
    n = np.sqrt((q * q).sum(-1, keepdims=True, dtype=np.float32))
    y = q / n
    dq = (d_rot - y * (y * d_rot).sum(-1, keepdims=True)) / n
    return dq.astype(np.float32), (d_opa * opa * (1 - opa)).astype(np.float32), (d_sca * sca).astype(np.float32)


def adam_step(p, g, m, v, step, lr, beta1=0.9, beta2=0.999, eps=1e-8):
    """One update; step is the 1-based count after the increment.  Returns (p, m, v) as float32."""
    f = np.float32
    p, g, m, v = (np.asarray(x, f) for x in (p, g, m, v))
    m = m + (g - m) * f(1 - beta1)
    v = v * f(beta2) + f(1 - beta2) * g * g
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = np.sqrt(v) / f(bc2 ** 0.5) + f(eps)
    p = p - f(lr / bc1) * (m / denom)
    return p.astype(f), m.astype(f), v.astype(f)
