"""Rigidity loss of the per-timestep training step on one fused kernel.

Drop-in names from train.py: NeighborInfo / ForegroundInfo (:42-52), create_neighbor_info (:166-183),
create_foreground_info (:226-247), calculate_rigidity_loss (:317-351).  The loss does not depend on the view,
but the reference evaluates it inside its per-view loop (train.py:373-377 in :402-412), i.e. 5 times per step with
~30 torch kernels each; here it is ONE launch that yields the loss and dloss/d(means, rotation_quaternions), evaluated
once per step (multiply by the number of views to reproduce the reference's summed loss).

The constant per-pair tables (neighbour indices, weights) are transposed to [K][F] int32 / float32 once per
NeighborInfo (coalesced for the kernel's thread-per-Gaussian mapping); ForegroundInfo.offsets_to_neighbors is a
[F,K,3] view of [K,3,F] memory for the same reason.  CUDA-only like the rest of gsb200.
"""
import weakref
from dataclasses import dataclass

import numpy as np
import torch

from . import _lib as L


@dataclass
class NeighborInfo:   # train.py:42-46
    weights: torch.Tensor
    indices: torch.Tensor


@dataclass
class ForegroundInfo:   # train.py:48-52
    inverted_rotations: torch.Tensor
    offsets_to_neighbors: torch.Tensor


def create_neighbor_info(gaussian_cloud_parameters, k=20):
    """train.py:166-183: k nearest foreground neighbours of every foreground Gaussian, weights exp(-2000 d^2)
    (scipy cKDTree in place of open3d, float64 distances)."""
    from .io import knn_squared_distances
    p = gaussian_cloud_parameters
    fg = p["segmentation_masks"][:, 0] > 0.5
    idx, sq = knn_squared_distances(p["means"][fg].detach().cpu().numpy(), k)
    dev = p["means"].device
    return NeighborInfo(indices=torch.tensor(idx).to(dev).long().contiguous(),
                        weights=torch.tensor(np.exp(-2000 * sq)).to(dev).float().contiguous())


def C_ptr(t):
    """Device pointer of a possibly non-contiguous view (L.ptr is for contiguous tensors by convention)."""
    import ctypes
    return ctypes.c_void_p(t.data_ptr()) if t.numel() else None


class _Prepared:
    """Kernel-side form of the constant tables of one (NeighborInfo, foreground mask) pair."""

    def __init__(self, neighbor_indices, weights, fg_mask):
        self.fg_index = torch.nonzero(fg_mask, as_tuple=False).squeeze(1).to(torch.int32).contiguous()
        self.F, self.K = neighbor_indices.shape
        if self.fg_index.numel() != self.F:
            raise RuntimeError(f"neighbour table has {self.F} rows but the mask selects {self.fg_index.numel()} Gaussians")
        self.nbr_t = neighbor_indices.t().to(torch.int32).contiguous()              # [K][F]
        self.w_t, self.w_src = None, None                                           # [K][F], set on first use


_cache = {}   # id(indices tensor) -> (weakref to it, key, _Prepared); entries die with the tensor


def _prepared(indices, weights, seg_masks):
    key = (indices.data_ptr(), seg_masks.data_ptr(), seg_masks.shape[0])
    slot = id(indices)
    hit = _cache.get(slot)
    if hit is None or hit[0]() is not indices or hit[1] != key:
        hit = (weakref.ref(indices, lambda _, slot=slot: _cache.pop(slot, None)), key,
               _Prepared(indices, None, (seg_masks[:, 0] > 0.5).detach()))
        _cache[slot] = hit
    prep = hit[2]
    if weights is not None and (prep.w_t is None or prep.w_src != weights.data_ptr()):
        prep.w_t, prep.w_src = weights.t().float().contiguous(), weights.data_ptr()
    return prep


def _offsets_storage(F, K, device):
    """[F,K,3] view over [K,3,F] memory: element (i,j,c) at j*3F + c*F + i."""
    return torch.empty((K, 3, F), dtype=torch.float32, device=device).permute(2, 0, 1)


def create_foreground_info(gaussian_cloud_parameters, initial_neighbor_indices):
    """train.py:226-247 in one launch."""
    p = gaussian_cloud_parameters
    means, quats = p["means"].detach().contiguous().float(), p["rotation_quaternions"].detach().contiguous().float()
    L.require_cuda(means, "means")
    prep = _prepared(initial_neighbor_indices, None, p["segmentation_masks"])
    F, K = prep.F, prep.K
    inv = torch.empty((F, 4), dtype=torch.float32, device=means.device)
    off = _offsets_storage(F, K, means.device)
    with L.device_guard(means.device):
        rc = L.lib().gsb_foreground_info(F, K, L.ptr(prep.fg_index), L.ptr(means), L.ptr(quats), L.ptr(prep.nbr_t), 1, F,
                                         L.ptr(inv), C_ptr(off), off.stride(0), off.stride(1), off.stride(2),
                                         L.stream_ptr())
    L.check(rc, "gsb_foreground_info")
    return ForegroundInfo(inverted_rotations=inv, offsets_to_neighbors=off)


class _Rigidity(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means, quats, prep, prev_inv, prev_off):
        m, q = means.contiguous().float(), quats.contiguous().float()
        need = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        loss = torch.empty((), dtype=torch.float64, device=m.device)
        dm = torch.zeros_like(m) if need else None
        dq = torch.zeros_like(q) if need else None
        F, K = prep.F, prep.K
        with L.device_guard(m.device):
            rc = L.lib().gsb_rigidity_loss(F, K, L.ptr(prep.fg_index), L.ptr(m), L.ptr(q), L.ptr(prev_inv),
                                           L.ptr(prep.nbr_t), 1, F, L.ptr(prep.w_t), 1, F, C_ptr(prev_off),
                                           prev_off.stride(0), prev_off.stride(1), prev_off.stride(2),
                                           C_ptr(loss), L.ptr(dm), L.ptr(dq), L.stream_ptr())
        L.check(rc, "gsb_rigidity_loss")
        if need:
            ctx.save_for_backward(dm, dq)
        return loss.float()

    @staticmethod
    def backward(ctx, g):
        dm, dq = ctx.saved_tensors
        return (dm * g if ctx.needs_input_grad[0] else None, dq * g if ctx.needs_input_grad[1] else None,
                None, None, None)


def calculate_rigidity_loss(gaussian_cloud_parameters, initial_neighbor_info, previous_timestep_foreground_info):
    """train.py:317-351: scalar loss, differentiable w.r.t. means and rotation_quaternions."""
    p = gaussian_cloud_parameters
    L.require_cuda(p["means"], "means")
    prep = _prepared(initial_neighbor_info.indices, initial_neighbor_info.weights, p["segmentation_masks"])
    prev = previous_timestep_foreground_info
    inv = prev.inverted_rotations.contiguous().float()
    off = prev.offsets_to_neighbors.float()
    if tuple(off.shape) != (prep.F, prep.K, 3) or tuple(inv.shape) != (prep.F, 4):
        raise RuntimeError("previous-timestep foreground info does not match the neighbour table")
    return _Rigidity.apply(p["means"], p["rotation_quaternions"], prep, inv, off)
