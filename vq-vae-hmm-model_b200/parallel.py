"""Batch sharding across the GPUs of one box: one process per GPU, sequences
partitioned, parameters replicated, NO communication inside any kernel, and exactly one NCCL
all-reduce per step over [flat gradients || loss sums] -- after a tiny all-reduce of the two
normalisers, unless the caller hands in the lengths of the whole batch (`global_lengths`), from which they are computed on the
host.  The sharded loss equals the un-sharded loss of the concatenated batch (the
reference divides the reconstruction term by the GLOBAL valid count, VQ_VAE_HMM_fixed.py:120)."""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import functional as Fn


def _global_norms(x, lengths_dev, norms_local, group, norms_global):
    """global normalisers: given by the caller (host-side, no collective) or the all-reduced sums of valid steps * C and of
    the batch sizes over the ranks"""
    if norms_global is not None:
        return norms_global.to(device=lengths_dev.device, dtype=torch.float64)
    C = x.shape[1]
    nv = lengths_dev.clamp(min=0, max=x.shape[2]).sum().to(torch.float64)
    cnt = torch.stack([nv * C, norms_local[1]])
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM, group=group)
    return torch.stack([cnt[0].clamp(min=1.0), cnt[1]])


class _ShardedElbo(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, x, u, lengths_dev, beta, norms_local, group, norms_global, *params):
        d = model._dims
        norms = _global_norms(x, lengths_dev, norms_local, group, norms_global)
        want = any(ctx.needs_input_grad)
        sums, grads = Fn.elbo_raw(d, params, x, u, lengths_dev, norms, beta, want)
        # one all-reduce: grads (fp32) followed by the sums split into hi/lo floats
        hi = sums.to(torch.float32)
        lo = (sums - hi.to(torch.float64)).to(torch.float32)
        parts = ([grads] if grads is not None else []) + [hi, lo]
        flat = torch.cat(parts)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        n = grads.numel() if grads is not None else 0
        gsums = flat[n:n + 4].to(torch.float64) + flat[n + 4:n + 8].to(torch.float64)
        ctx.has_grads = grads is not None
        ctx.shapes = [tuple(p.shape) for p in params]
        ctx.offsets = Fn.param_offsets(d)
        ctx.save_for_backward(flat[:n] if n else gsums)
        return Fn.loss_from_sums(gsums, norms, beta)

    @staticmethod
    def backward(ctx, gout):
        (flat,) = ctx.saved_tensors
        if not ctx.has_grads:
            raise Fn.VhmmError("ELBO gradients were not computed")
        flat = flat * gout.to(flat.dtype)
        outs = [flat[ctx.offsets[i]:ctx.offsets[i + 1]].view(s) for i, s in enumerate(ctx.shapes)]
        return (None, None, None, None, None, None, None, None, *outs)


def sharded_loss_and_flat_grads(model, x, u, lengths_dev, beta, norms_local, group=None, norms_global=None):
    """The exchange of _ShardedElbo without autograd: (global loss, GLOBAL flat gradient buffer)."""
    d = model._dims
    norms = _global_norms(x, lengths_dev, norms_local, group, norms_global)
    sums, grads = Fn.elbo_raw(d, model._params(), x, u, lengths_dev, norms, beta, True)
    hi = sums.to(torch.float32)
    lo = (sums - hi.to(torch.float64)).to(torch.float32)
    flat = torch.cat([grads, hi, lo])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    n = grads.numel()
    gsums = flat[n:n + 4].to(torch.float64) + flat[n + 4:n + 8].to(torch.float64)
    return Fn.loss_from_sums(gsums, norms, beta), flat[:n]


def sharded_elbo(model, x, u, lengths_dev, beta, norms_local, group=None, norms_global=None):
    """Global loss (identical on every rank) whose backward leaves the GLOBAL gradient in .grad."""
    return _ShardedElbo.apply(model, x, u, lengths_dev, beta, norms_local, group, norms_global, *model._params())


def global_norms_from_lengths(global_lengths, C, T, device):
    """The two normalisers of the WHOLE batch (valid steps * C clamped at 1, sequences: VQ_VAE_HMM_fixed.py:120, :131) from
    the lengths of every rank's sequences, computed on the host -- a caller that shards a batch it holds (shard_batch) knows
    them before the split, and the step then needs ONE collective instead of two."""
    gl = torch.as_tensor(global_lengths).to(torch.int64).cpu()
    nv = int(gl.clamp(min=0, max=T).sum())
    return torch.tensor([float(max(nv * C, 1)), float(gl.numel())], dtype=torch.float64, device=device)


def shard_batch(x, u, lengths, rank, world):
    """Contiguous split of the batch dimension: rank g gets [g*B/G, (g+1)*B/G)."""
    B = x.shape[0]
    lo, hi = (B * rank) // world, (B * (rank + 1)) // world
    return x[lo:hi], u[lo:hi], lengths[lo:hi]


def bind_to_gpu_numa_node(device_index):
    """Pin this process to the CPU cores of the NUMA node its GPU hangs off (sysfs), BEFORE the pinned staging buffers are
    allocated: first-touch then places them on the local node, and the host->device streams of the ranks of one box stop
    crossing the socket interconnect.  Returns the node, or None when it cannot be determined (single
    node, containers without sysfs, no permission) -- never raises."""
    import os
    try:
        import torch
        bus = torch.cuda.get_device_properties(device_index).pci_bus_id \
            if hasattr(torch.cuda.get_device_properties(device_index), "pci_bus_id") else None
        dom = getattr(torch.cuda.get_device_properties(device_index), "pci_domain_id", 0)
        dev = getattr(torch.cuda.get_device_properties(device_index), "pci_device_id", 0)
        if bus is None:
            return None
        path = f"/sys/bus/pci/devices/{dom:04x}:{bus:02x}:{dev:02x}.0/numa_node"
        node = int(open(path).read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = os.sched_getaffinity(0) & cpus
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return node
    except Exception:
        return None


def enable_sharding(model, group=None):
    """After this, model.compute_loss() returns the global loss / leaves global gradients."""
    model.shard_group = group if group is not None else dist.group.WORLD
    return model
