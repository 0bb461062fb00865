"""TEST INFRASTRUCTURE ONLY (imported by tests/ and nothing else): numpy restatement of the reference's rigidity loss
and its gradient, the checker for gsb_rigidity_loss (csrc/rigidity.cu).

Follows /root/reference/train.py:311-351 (calculate_rigidity_loss), :301-308 (quat_mult), :314-315
(weighted_l2_loss_v2) and /root/reference/external.py:27-47 (build_rotation):

    fg           = segmentation_masks[:, 0] > 0.5
    q            = normalize(rotation_quaternions[fg])                 (F.normalize, eps 1e-12)
    rel          = quat_mult(q, previous.inverted_rotations)           (Hamilton product, (w, x, y, z))
    R            = build_rotation(rel)                                 (normalises rel again, no eps)
    off[i, j]    = means[fg][nbr[i, j]] - means[fg][i]
    rot[i, j]    = R[i]^T off[i, j]
    loss         = mean_ij sqrt(|rot - previous.offsets_to_neighbors|^2 * w[i, j] + 1e-20)

Pinned to the reference's own code run on CPU (tests/golden/make_golden_rigidity.py -> tests/golden/rigidity_*.npz,
tests/test_rigidity.py).  Everything is evaluated in float64 from float32 inputs unless dtype is given.
"""
import numpy as np


def quat_mult(q1, q2):
    w1, x1, y1, z1 = q1.T
    w2, x2, y2, z2 = q2.T
    return np.stack([w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2,
                     w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                     w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2], axis=-1)


def build_rotation(q):
    q = q / np.sqrt((q * q).sum(-1))[:, None]
    r, x, y, z = q.T
    R = np.empty((q.shape[0], 3, 3), q.dtype)
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r * z); R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y); R[:, 2, 1] = 2 * (y * z + r * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R, q


def rigidity_loss(means, quats, fg_index, nbr, weights, prev_inv_rot, prev_offsets, dtype=np.float64):
    """means [P,3], quats [P,4] raw, fg_index [F] rows of the foreground Gaussians, nbr [F,K] (foreground space),
    weights [F,K], prev_inv_rot [F,4], prev_offsets [F,K,3].
    Returns (loss, d_means [P,3], d_quats [P,4]) with d = dloss/d(raw input)."""
    means, quats = np.asarray(means, dtype), np.asarray(quats, dtype)
    w, p, prev = np.asarray(weights, dtype), np.asarray(prev_inv_rot, dtype), np.asarray(prev_offsets, dtype)
    fg_index, nbr = np.asarray(fg_index), np.asarray(nbr)
    F, K = nbr.shape
    m = means[fg_index]
    q_raw = quats[fg_index]
    n = np.maximum(np.sqrt((q_raw * q_raw).sum(-1)), 1e-12)
    q = q_raw / n[:, None]
    rel = quat_mult(q, p)
    rel_n = np.sqrt((rel * rel).sum(-1))
    R, u = build_rotation(rel)
    off = m[nbr] - m[:, None]                                  # [F,K,3]
    rot = np.einsum("fba,fkb->fka", R, off)                    # R^T off
    d = rot - prev
    s = (d * d).sum(-1) * w + 1e-20
    root = np.sqrt(s)
    loss = root.mean()
    # ---- gradient ----
    g = (w / root)[..., None] * d / (F * K)                    # dloss/drot
    d_off = np.einsum("fba,fka->fkb", R, g)                    # R g
    dR = np.einsum("fkb,fka->fba", off, g)                     # dR[b][a] = sum_k off_b g_a
    d_m = np.zeros_like(m)
    np.add.at(d_m, nbr.reshape(-1), d_off.reshape(-1, 3))
    d_m -= d_off.sum(1)
    r, x, y, z = u.T
    D = lambda b, a: dR[:, b, a]
    du = np.stack([
        2 * (-z * D(0, 1) + y * D(0, 2) + z * D(1, 0) - x * D(1, 2) - y * D(2, 0) + x * D(2, 1)),
        2 * (y * D(0, 1) + z * D(0, 2) + y * D(1, 0) - 2 * x * D(1, 1) - r * D(1, 2) + z * D(2, 0) + r * D(2, 1)
             - 2 * x * D(2, 2)),
        2 * (-2 * y * D(0, 0) + x * D(0, 1) + r * D(0, 2) + x * D(1, 0) + z * D(1, 2) - r * D(2, 0) + z * D(2, 1)
             - 2 * y * D(2, 2)),
        2 * (-2 * z * D(0, 0) - r * D(0, 1) + x * D(0, 2) + r * D(1, 0) - 2 * z * D(1, 1) + y * D(1, 2) + x * D(2, 0)
             + y * D(2, 1))], axis=-1)
    d_rel = (du - u * (u * du).sum(-1, keepdims=True)) / rel_n[:, None]
    dw, dx, dy, dz = d_rel.T
    w2, x2, y2, z2 = p.T
    dq = np.stack([dw * w2 + dx * x2 + dy * y2 + dz * z2,
                   -dw * x2 + dx * w2 - dy * z2 + dz * y2,
                   -dw * y2 + dx * z2 + dy * w2 - dz * x2,
                   -dw * z2 - dx * y2 + dy * x2 + dz * w2], axis=-1)
    dq_raw = (dq - q * (q * dq).sum(-1, keepdims=True)) / n[:, None]
    d_means, d_quats = np.zeros_like(means), np.zeros_like(quats)
    np.add.at(d_means, fg_index, d_m)
    np.add.at(d_quats, fg_index, dq_raw)
    return loss, d_means, d_quats


def foreground_info(means, quats, fg_index, nbr, dtype=np.float64):
    """create_foreground_info (train.py:226-247): inverted normalised rotations and offsets to neighbours of the
    previous timestep."""
    q = np.asarray(quats, dtype)[fg_index]
    q = q / np.maximum(np.sqrt((q * q).sum(-1)), 1e-12)[:, None]
    q[:, 1:] *= -1
    m = np.asarray(means, dtype)[fg_index]
    return q, m[nbr] - m[:, None]
