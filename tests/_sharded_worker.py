"""Worker of tests/test_sharded.py: one rank of a world_size-2 gloo job (MASTER_ADDR 127.0.0.1).

mode "cpu":  the model stays on the CPU and functional.elbo_raw -- the ONE call into libvhmm.so on this
             path -- is replaced, IN THIS TEST PROCESS ONLY, by the float64 oracle (un-normalised sums + the
             gradient under the GLOBAL normalisers, exactly the contract of vhmm_elbo); everything else is
             the product's own host logic: parallel._ShardedElbo, the normaliser exchange, the hi/lo split
             of the sums, the single all-reduce, autograd scaling.
mode "gpu":  both ranks share cuda:0 and run the real kernels; gloo carries the (CUDA) tensors.
Each rank saves {loss, grads...} to an .npz; the test compares them with the un-sharded oracle.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_batch():
    from oracle import elbo as O
    x, u, L = O.synth_batch(12, 5, 4, 48, seed=77, ragged=True, min_len=3)
    L[3] = 1
    L[8] = 0       # an empty sequence on the second shard
    keep = (torch.arange(48)[None, :] < L[:, None]).unsqueeze(1)
    return x * keep, u * keep, L


def shard(rank):   # uneven on purpose: 5 + 7 sequences with different valid counts
    return (0, 5) if rank == 0 else (5, 12)


def main():
    mode, rank, world, port, out = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=port, RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import vhmm_b200 as V
    from vhmm_b200 import functional as Fn
    from oracle import elbo as O
    if mode == "cpu":
        def oracle_elbo_raw(d, params, x, u, lengths_dev, norms, beta, want_grads):
            with torch.enable_grad():      # (autograd.Function.forward runs with grad mode off)
                p = {k: t.detach().double().requires_grad_(True) for k, t in zip(O.PARAM_ORDER, params)}
                B, C, T = x.shape
                recon, prior_loss, ent = O.loss_terms(p, x.double(), u.double(), lengths_dev, d.U)
                nv = float(lengths_dev.clamp(min=0, max=T).sum())
                s0, s1, s2 = recon * max(nv * C, 1.0), -prior_loss * B, ent * B       # undo the LOCAL normalisers
                sums = torch.stack([s0, s1, s2, torch.zeros((), dtype=torch.float64)]).detach()
                grads = None
                if want_grads:
                    loss = s0 / norms[0] + beta * (-s1 / norms[1] - s2 / norms[1])    # GLOBAL normalisers (vhmm_elbo contract)
                    g = torch.autograd.grad(loss, [p[k] for k in O.PARAM_ORDER], allow_unused=True)
                    grads = torch.cat([(gi if gi is not None else torch.zeros_like(p[k])).flatten()
                                       for gi, k in zip(g, O.PARAM_ORDER)]).float()
            return sums, grads
        Fn.elbo_raw = oracle_elbo_raw
        dev = torch.device("cpu")
    else:
        dev = torch.device("cuda", 0)
        torch.cuda.set_device(0)
    p = O.init_params(5, 64, 3, 32, 4, 128, seed=3)
    m = V.VAE_HMM(5, 64, 3, 32, u_dim=4)
    m.load_state_dict(p)
    m = m.to(dev)
    V.parallel.enable_sharding(m)
    x, u, L = make_batch()
    lo, hi = shard(rank)
    res = {}
    for tag, beta in (("a", 1.0), ("b", 0.37)):
        m.zero_grad(set_to_none=True)
        loss = m.compute_loss(x[lo:hi].to(dev), u[lo:hi].to(dev), L[lo:hi], beta)
        loss.backward()
        res[f"loss_{tag}"] = loss.detach().double().cpu().numpy()
        for k, t in m.named_parameters():
            res[f"grad_{tag}:{k}"] = t.grad.detach().double().cpu().numpy()
    # the same step with the global normalisers computed on the host from the lengths of the whole batch (one collective)
    m.zero_grad(set_to_none=True)
    loss = m.compute_loss(x[lo:hi].to(dev), u[lo:hi].to(dev), L[lo:hi], 0.37, global_lengths=L)
    loss.backward()
    res["loss_c"] = loss.detach().double().cpu().numpy()
    for k, t in m.named_parameters():
        res[f"grad_c:{k}"] = t.grad.detach().double().cpu().numpy()
    l2, flat = m.loss_and_flat_grads(x[lo:hi].to(dev), u[lo:hi].to(dev), L[lo:hi], 0.37)
    res["loss_flat"] = l2.double().cpu().numpy()
    res["flat"] = flat.double().cpu().numpy()
    np.savez(out, **res)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
