"""DeformationNetwork forward on the tcgen05 dense layers.

Mirrors train.py: ResidualBlock / DeformationNetwork (:57-110), PositionalEncoding (:113-127),
normalize_and_encode_means_and_rotations (:186-205), update_gaussian_cloud_parameters (:268-306).  The module keeps
the reference's parameter names (fc_in, residual_blocks.<i>.fc1 / bn1 / fc2 / bn2, fc_out), so state_dicts are
interchangeable with the reference's nn.Module in both directions.

Per layer: ONE tcgen05 GEMM whose epilogue also accumulates the BatchNorm batch statistics, then ONE pass that
normalises, adds the residual and applies GELU.  BatchNorm always uses batch statistics and updates its running
buffers, because the reference never leaves training mode.  `precise=True` (the default) selects the 3xTF32 split
(FP32-level accuracy: the reference never enables allow_tf32, so its nn.Linear layers are FP32 products);
`precise=False` is an opt-in fast mode in plain TF32 (what torch computes with allow_tf32), forward and backward.

Training: when gradients are enabled the forward runs inside ONE autograd node (_DeformFn) that keeps the per-block
activations (x, z1, a, z2 and the BatchNorm sums) and whose backward is built from the same pieces: dX = dY W on the
forward GEMM with the transposed weight, dW = dY^T X on the MN-major tcgen05 weight-gradient kernel, and the two-pass
BatchNorm / GELU adjoint (gsb_bn_gelu_backward).  With precise=True the input gradients run on the 3xTF32 layer and
the weight gradients on the 3xTF32 form of the weight-gradient kernel (both streamed operands split into hi / lo where
they land in shared memory: gsb_linear_wgrad_tf32x3); with precise=False both run in plain TF32.  The 7-wide output
layer's two backward products (7 is not a tensor-core tile) are FP32 streaming kernels (gsb_linear_thin_dx /
gsb_linear_thin_wgrad).  No torch / cuBLAS matmul is left in the network.

Row sharding (multi-GPU; the reference has no distributed code): the network is a per-Gaussian function except for the
BatchNorm batch statistics, so with `net.set_row_shard(group)` every rank runs it on its contiguous slice of the
Gaussians only: the column sums behind every BatchNorm (forward: sum z, sum z^2; backward: sum g, sum g xhat -- 2 x 256
doubles each) are added over the ranks, the [N, 7] output is all-gathered, and the parameter gradients (one flat
buffer) are summed at the end of the backward.  The result equals the replicated network up to summation order, and
replicas stay bit-identical because every rank receives the same reduced numbers.  update_gaussian_cloud_parameters
uses it automatically when the network carries a shard.
"""
import torch
from torch import nn

import torch.distributed as dist

from . import _lib as L
from .linear import linear, linear_wgrad, thin_backward


class RowShard:
    """This rank's contiguous slice of the Gaussians under row sharding, and the two collectives the network needs.
    Works over NCCL (device tensors) and over gloo (tests: device tensors are staged through the host)."""

    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._via_host = dist.is_initialized() and dist.get_backend(group) == "gloo"

    def bounds(self, n, rank=None):
        """Rows [r0, r1) of rank `rank` (default: this rank): equal chunks of ceil(n / world) rows, the last ones shorter."""
        rank = self.rank if rank is None else rank
        chunk = (n + self.world - 1) // self.world
        return min(rank * chunk, n), min((rank + 1) * chunk, n)

    def all_reduce(self, t):
        if self.world > 1:
            if self._via_host and t.is_cuda:
                h = t.cpu()
                dist.all_reduce(h, group=self.group)
                t.copy_(h)
            else:
                dist.all_reduce(t, group=self.group)
        return t

    def all_gather_rows(self, local, n):
        """local [r1 - r0, k] of every rank -> [n, k] (rank order)."""
        if self.world == 1:
            return local
        chunk = (n + self.world - 1) // self.world
        pad = torch.zeros((chunk, local.shape[1]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        if self._via_host and pad.is_cuda:
            parts = [torch.empty(pad.shape, dtype=pad.dtype) for _ in range(self.world)]
            dist.all_gather(parts, pad.cpu(), group=self.group)
            out = torch.cat(parts, dim=0).to(local.device)
        else:
            out = torch.empty((self.world * chunk, local.shape[1]), dtype=local.dtype, device=local.device)
            dist.all_gather_into_tensor(out, pad, group=self.group)
        return out[:n]


class _GatherRows(torch.autograd.Function):
    """All-gather of the ranks' row slices.  Every rank goes on to evaluate the SAME loss on the gathered tensor (the
    image term's gradient is already summed over the ranks by the step's exchange, the rigidity term is replicated),
    so the adjoint is this rank's slice of the incoming gradient, not a reduce-scatter."""

    @staticmethod
    def forward(ctx, local, shard, n):
        ctx.bounds = shard.bounds(n)
        return shard.all_gather_rows(local, n)

    @staticmethod
    def backward(ctx, g):
        r0, r1 = ctx.bounds
        return g[r0:r1].contiguous(), None, None


def positional_encoding(x, frequency_count, lo=None, span=None, out=None, col0=0):
    """PositionalEncoding(frequency_count)(x) (train.py:113-127) for x [N, D]; optional min/span normalisation first.
    Writes into out[:, col0 : col0 + 2 F D] when `out` is given."""
    L.require_cuda(x, "x")
    x = x.detach().float()
    if x.stride(-1) != 1:
        x = x.contiguous()
    n, d = x.shape
    if out is None:
        out = torch.empty((n, 2 * frequency_count * d), dtype=torch.float32, device=x.device)
        col0 = 0
    with L.device_guard(x.device):
        rc = L.lib().gsb_positional_encoding(n, d, frequency_count, L.ptr(x), x.stride(0), L.ptr(lo), L.ptr(span), L.ptr(out),
                                             out.stride(0), col0, L.stream_ptr())
    L.check(rc, "gsb_positional_encoding")
    return out


def _min_span(t):
    lo = t.min(dim=0).values
    return lo.contiguous(), (t - lo).max(dim=0).values.contiguous()


def normalize_and_encode_means_and_rotations(gaussian_cloud_parameters, out=None, col0=0):
    """train.py:186-205: [N, 92] = PE10(normalised means) | PE4(normalised rotations)."""
    means = gaussian_cloud_parameters["means"].detach().float()
    rots = gaussian_cloud_parameters["rotation_quaternions"].detach().float()
    n = means.shape[0]
    if out is None:
        out = torch.empty((n, 92), dtype=torch.float32, device=means.device)
        col0 = 0
    positional_encoding(means, 10, *_min_span(means), out=out, col0=col0)
    positional_encoding(rots, 4, *_min_span(rots), out=out, col0=col0 + 60)
    return out


class ResidualBlock(nn.Module):   # parameter container with the reference's names (train.py:57-78)
    def __init__(self, dimension):
        super().__init__()
        self.fc1 = nn.Linear(dimension, dimension, bias=False)
        self.bn1 = nn.BatchNorm1d(dimension)
        self.fc2 = nn.Linear(dimension, dimension, bias=False)
        self.bn2 = nn.BatchNorm1d(dimension)


class DeformationNetwork(nn.Module):
    def __init__(self, hidden_dimension, residual_block_count, precise=True):
        super().__init__()
        if hidden_dimension % 4 or hidden_dimension > 256:
            raise ValueError("hidden_dimension must be a multiple of 4 and at most 256 on the tcgen05 path")
        self.fc_in = nn.Linear(192, hidden_dimension)
        self.residual_blocks = nn.Sequential(*(ResidualBlock(hidden_dimension) for _ in range(residual_block_count)))
        self.fc_out = nn.Linear(hidden_dimension, 7)
        self.precise = precise
        self.shard = None          # RowShard: the network runs on this rank's slice of the rows (set_row_shard)
        self._batch_rows = {}

    def set_row_shard(self, group=None):
        """Row-sharded mode over the ranks of `group` (see the module docstring); world size 1 leaves the network as it is."""
        shard = RowShard(group)
        self.shard = shard if shard.world > 1 else None
        self._batch_rows = {}
        return self

    def _rows_of_batch(self, rows, device):
        """Rows of the whole batch = sum of the ranks' rows (one tiny all-reduce per distinct local row count)."""
        if self.shard is None:
            return rows
        if rows not in self._batch_rows:
            t = torch.tensor([rows], dtype=torch.int64, device=device)
            self._batch_rows[rows] = int(self.shard.all_reduce(t).item())
        return self._batch_rows[rows]

    def _bn_act(self, z, stats, bn, residual, gelu=True):
        rows, n = z.shape
        y = torch.empty_like(z)
        batch_rows = self._rows_of_batch(rows, z.device)
        if self.shard is not None:
            self.shard.all_reduce(stats)     # column sums of the whole batch
        with L.device_guard(z.device):
            rc = L.lib().gsb_bn_gelu(rows, batch_rows, n, L.ptr(z), z.stride(0), L.ptr(stats), L.ptr(bn.weight.detach()),
                                     L.ptr(bn.bias.detach()), float(bn.eps), L.ptr(residual),
                                     0 if residual is None else residual.stride(0), L.ptr(y), y.stride(0),
                                     L.ptr(bn.running_mean), L.ptr(bn.running_var),
                                     float(bn.momentum if bn.momentum is not None else 0.1), 1 if gelu else 0, L.stream_ptr())
        L.check(rc, "gsb_bn_gelu")
        bn.num_batches_tracked += 1
        return y

    def _bn_bwd(self, dy, z, stats, bn, residual, gelu=True):
        """-> (dz, g, d_gamma, d_beta): g = dy * act'(.) is also the residual branch's gradient."""
        rows, n = z.shape
        g = torch.empty_like(z)
        dz = torch.empty_like(z) if residual is not None else g
        sums = torch.empty((2, n), dtype=torch.float64, device=z.device)
        dgam, dbet = torch.empty(n, device=z.device), torch.empty(n, device=z.device)
        dy = dy.contiguous()
        batch_rows = self._rows_of_batch(rows, z.device)

        def run(phase):
            with L.device_guard(z.device):
                rc = L.lib().gsb_bn_gelu_backward(rows, batch_rows, phase, n, L.ptr(dy), dy.stride(0), L.ptr(z), z.stride(0),
                                                  L.ptr(stats), L.ptr(bn.weight.detach()), L.ptr(bn.bias.detach()),
                                                  float(bn.eps), L.ptr(residual), 0 if residual is None else residual.stride(0),
                                                  1 if gelu else 0, L.ptr(g), g.stride(0), L.ptr(dz), dz.stride(0), L.ptr(sums),
                                                  L.ptr(dgam), L.ptr(dbet), L.stream_ptr())
            L.check(rc, "gsb_bn_gelu_backward")

        if self.shard is None:
            run(0)
        else:   # the two column sums of the adjoint cover the whole batch: add the ranks' sums between the passes
            run(1)
            self.shard.all_reduce(sums)
            run(2)
        return dz, g, dgam, dbet

    def _forward_impl(self, x, keep):
        """Shared by inference and training; keep = list that receives the per-block activations (or None)."""
        h = self.fc_in.out_features
        out = linear(x, self.fc_in.weight, self.fc_in.bias, precise=self.precise)
        for block in self.residual_blocks:
            identity = out
            st1 = torch.zeros((2, h), dtype=torch.float64, device=x.device)
            z1 = linear(identity, block.fc1.weight, None, precise=self.precise, col_stats=st1)
            a = self._bn_act(z1, st1, block.bn1, None)
            st2 = torch.zeros((2, h), dtype=torch.float64, device=x.device)
            z2 = linear(a, block.fc2.weight, None, precise=self.precise, col_stats=st2)
            out = self._bn_act(z2, st2, block.bn2, identity)
            if keep is not None:
                keep.append((identity, z1, st1, a, z2, st2))
        delta = linear(out, self.fc_out.weight, self.fc_out.bias, precise=self.precise)
        return delta, out

    def _parameter_list(self):
        ps = [self.fc_in.weight, self.fc_in.bias]
        for b in self.residual_blocks:
            ps += [b.fc1.weight, b.bn1.weight, b.bn1.bias, b.fc2.weight, b.bn2.weight, b.bn2.bias]
        return ps + [self.fc_out.weight, self.fc_out.bias]

    def forward(self, initial_means_and_rotations, encoded_normalized_initial_means_and_rotations=None,
                encoded_normalized_previous_means_and_rotations=None, encoded_progress=None, network_input=None):
        """train.py:88-110.  Pass either the three encoded blocks (concatenated here, like the reference) or the
        already assembled [N, 192] `network_input`."""
        if network_input is None:
            network_input = torch.cat((encoded_normalized_initial_means_and_rotations,
                                       encoded_normalized_previous_means_and_rotations, encoded_progress), dim=1)
        x = network_input.detach().float().contiguous()
        params = self._parameter_list()
        if torch.is_grad_enabled() and any(p.requires_grad for p in params):
            return _DeformFn.apply(self, x, initial_means_and_rotations, *params)
        with torch.no_grad():
            delta, _ = self._forward_impl(x, None)
            return delta + initial_means_and_rotations


class _DeformFn(torch.autograd.Function):
    """The whole network as one autograd node (inputs: the [N,192] encoding (no gradient), the [N,7] initial means and
    rotations, the parameters in DeformationNetwork._parameter_list order)."""

    @staticmethod
    def forward(ctx, net, x, init, *params):
        keep = []
        delta, last = net._forward_impl(x, keep)
        ctx.net, ctx.x, ctx.keep, ctx.last = net, x, keep, last
        return delta + init

    @staticmethod
    def backward(ctx, d_out):
        net, x, keep, last = ctx.net, ctx.x, ctx.keep, ctx.last
        d_out = d_out.contiguous().float()
        grads = {}
        # precise (the default, = the reference's FP32 matmuls): input gradients on the 3xTF32 tcgen05 layer, weight
        # gradients as FP32 products; precise=False: both in plain TF32 on the tcgen05 kernels (what torch computes
        # with allow_tf32) -- an opt-in fast mode
        precise = net.precise

        def dx(g, w, residual=None):
            return linear(g, w.detach().t().contiguous(), None, precise=precise, residual=residual)

        def dw(g, a):
            return linear_wgrad(g, a, precise=precise)

        # output layer (7 columns, not a tensor-core tile): FP32 streaming kernels
        dh, grads[net.fc_out.weight], grads[net.fc_out.bias] = thin_backward(d_out, net.fc_out.weight, last)
        for block, (identity, z1, st1, a, z2, st2) in zip(reversed(list(net.residual_blocks)), reversed(keep)):
            dz2, g2, dgam2, dbet2 = net._bn_bwd(dh, z2, st2, block.bn2, identity)
            grads[block.bn2.weight], grads[block.bn2.bias] = dgam2, dbet2
            grads[block.fc2.weight] = dw(dz2, a)
            da = dx(dz2, block.fc2.weight)
            dz1, _, dgam1, dbet1 = net._bn_bwd(da, z1, st1, block.bn1, None)
            grads[block.bn1.weight], grads[block.bn1.bias] = dgam1, dbet1
            grads[block.fc1.weight] = dw(dz1, identity)
            dh = dx(dz1, block.fc1.weight, residual=g2)   # + the residual branch's gradient, added in the GEMM epilogue
        grads[net.fc_in.weight] = dw(dh, x)
        grads[net.fc_in.bias] = dh.sum(0)
        ctx.keep = ctx.last = ctx.x = None
        plist = net._parameter_list()
        if net.shard is not None:
            # row-sharded: the dense layers' gradients are sums over this rank's rows -> ONE all-reduce of a flat buffer
            # (the BatchNorm gradients come from the already exchanged column sums)
            dense = [p for p in plist if p.dim() == 2 or p is net.fc_in.bias or p is net.fc_out.bias]
            flat = torch.cat([grads[p].reshape(-1) for p in dense])
            net.shard.all_reduce(flat)
            o = 0
            for p in dense:
                grads[p] = flat[o:o + p.numel()].view_as(p)
                o += p.numel()
        return (None, None, d_out) + tuple(grads[p] if p.requires_grad else None for p in plist)


def update_gaussian_cloud_parameters(deformation_network, initial_gaussian_cloud_parameters,
                                     encoded_normalized_initial_means_and_rotations,
                                     encoded_normalized_previous_means_and_rotations, timestep, timestep_count):
    """train.py:268-306: means += 0.01 delta[:, :3], rotation_quaternions += 0.01 delta[:, 3:] (differentiable w.r.t. the
    network's parameters).  The [N, 192] input is assembled without the torch.cat: the progress encoding is written
    in place."""
    p = initial_gaussian_cloud_parameters
    n, dev = p["means"].shape[0], p["means"].device
    shard = getattr(deformation_network, "shard", None)
    r0, r1 = shard.bounds(n) if shard is not None else (0, n)   # row-sharded network: this rank's Gaussians only
    x = torch.empty((r1 - r0, 192), dtype=torch.float32, device=dev)
    x[:, :92] = encoded_normalized_initial_means_and_rotations[r0:r1]
    x[:, 92:184] = encoded_normalized_previous_means_and_rotations[r0:r1]
    progress = torch.full((r1 - r0, 1), timestep / timestep_count, dtype=torch.float32, device=dev)
    positional_encoding(progress, 4, out=x, col0=184)
    delta = deformation_network(torch.cat((p["means"][r0:r1], p["rotation_quaternions"][r0:r1]), dim=1).detach(),
                                network_input=x)
    if shard is not None:
        delta = _GatherRows.apply(delta, shard, n)
    updated = dict(p)
    updated["means"] = p["means"].detach() + delta[:, :3] * 0.01
    updated["rotation_quaternions"] = p["rotation_quaternions"].detach() + delta[:, 3:] * 0.01
    return updated
