"""DeformationNetwork forward on the tcgen05 dense layers.

Mirrors train.py: ResidualBlock / DeformationNetwork (:57-110), PositionalEncoding (:113-127),
normalize_and_encode_means_and_rotations (:186-205), update_gaussian_cloud_parameters (:268-306).  The module keeps
the reference's parameter names (fc_in, residual_blocks.<i>.fc1 / bn1 / fc2 / bn2, fc_out), so state_dicts are
interchangeable with the reference's nn.Module in both directions.

Per layer: ONE tcgen05 GEMM whose epilogue also accumulates the BatchNorm batch statistics, then ONE pass that
normalises, adds the residual and applies GELU.  BatchNorm always uses batch statistics and updates its running
buffers, because the reference never leaves training mode.  `precise=True` selects the 3xTF32 split (FP32-level
accuracy); the default is plain TF32 (what torch does with allow_tf32).

FORWARD ONLY in this round: outputs carry no autograd graph.  Training the network needs the backward GEMMs
(dX = dY W, dW = dY^T X) and the BatchNorm / GELU adjoints, which are the next step of this row.
"""
import torch
from torch import nn

from . import _lib as L
from .linear import linear


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
    def __init__(self, hidden_dimension, residual_block_count, precise=False):
        super().__init__()
        if hidden_dimension % 4 or hidden_dimension > 256:
            raise ValueError("hidden_dimension must be a multiple of 4 and at most 256 on the tcgen05 path")
        self.fc_in = nn.Linear(192, hidden_dimension)
        self.residual_blocks = nn.Sequential(*(ResidualBlock(hidden_dimension) for _ in range(residual_block_count)))
        self.fc_out = nn.Linear(hidden_dimension, 7)
        self.precise = precise

    def _bn_act(self, z, stats, bn, residual, gelu=True):
        rows, n = z.shape
        y = torch.empty_like(z)
        with L.device_guard(z.device):
            rc = L.lib().gsb_bn_gelu(rows, n, L.ptr(z), z.stride(0), L.ptr(stats), L.ptr(bn.weight.detach()),
                                     L.ptr(bn.bias.detach()), float(bn.eps), L.ptr(residual),
                                     0 if residual is None else residual.stride(0), L.ptr(y), y.stride(0),
                                     L.ptr(bn.running_mean), L.ptr(bn.running_var),
                                     float(bn.momentum if bn.momentum is not None else 0.1), 1 if gelu else 0, L.stream_ptr())
        L.check(rc, "gsb_bn_gelu")
        bn.num_batches_tracked += 1
        return y

    @torch.no_grad()
    def forward(self, initial_means_and_rotations, encoded_normalized_initial_means_and_rotations=None,
                encoded_normalized_previous_means_and_rotations=None, encoded_progress=None, network_input=None):
        """train.py:88-110.  Pass either the three encoded blocks (concatenated here, like the reference) or the
        already assembled [N, 192] `network_input`."""
        if network_input is None:
            network_input = torch.cat((encoded_normalized_initial_means_and_rotations,
                                       encoded_normalized_previous_means_and_rotations, encoded_progress), dim=1)
        x = network_input.float().contiguous()
        h = self.fc_in.out_features
        out = linear(x, self.fc_in.weight, self.fc_in.bias, precise=self.precise)
        for block in self.residual_blocks:
            identity = out
            stats = torch.zeros((2, h), dtype=torch.float64, device=x.device)
            z = linear(identity, block.fc1.weight, None, precise=self.precise, col_stats=stats)
            a = self._bn_act(z, stats, block.bn1, None)
            stats = torch.zeros((2, h), dtype=torch.float64, device=x.device)
            z = linear(a, block.fc2.weight, None, precise=self.precise, col_stats=stats)
            out = self._bn_act(z, stats, block.bn2, identity)
        delta = linear(out, self.fc_out.weight, self.fc_out.bias, precise=self.precise)
        return delta + initial_means_and_rotations


def update_gaussian_cloud_parameters(deformation_network, initial_gaussian_cloud_parameters,
                                     encoded_normalized_initial_means_and_rotations,
                                     encoded_normalized_previous_means_and_rotations, timestep, timestep_count):
    """train.py:268-306 (forward only): means += 0.01 delta[:, :3], rotation_quaternions += 0.01 delta[:, 3:].
    The [N, 192] input is assembled without the torch.cat: the progress encoding is written in place."""
    p = initial_gaussian_cloud_parameters
    n, dev = p["means"].shape[0], p["means"].device
    x = torch.empty((n, 192), dtype=torch.float32, device=dev)
    x[:, :92] = encoded_normalized_initial_means_and_rotations
    x[:, 92:184] = encoded_normalized_previous_means_and_rotations
    progress = torch.full((n, 1), timestep / timestep_count, dtype=torch.float32, device=dev)
    positional_encoding(progress, 4, out=x, col0=184)
    delta = deformation_network(torch.cat((p["means"], p["rotation_quaternions"]), dim=1).detach(), network_input=x)
    updated = dict(p)
    updated["means"] = p["means"].detach() + delta[:, :3] * 0.01
    updated["rotation_quaternions"] = p["rotation_quaternions"].detach() + delta[:, 3:] * 0.01
    return updated
