"""Multi-GPU ensembles: one process per GPU, trajectories sharded with NO per-step collective.

The reference has no ensemble driver (its idiom is a serial loop over `taylorinteg`, docs/src/root_finding.md:121-132);
trajectories never interact, so the only exchanges are
  * one all-gather of the per-trajectory results (`xfinal`, `tfinal`, `nsteps`, `status`, optionally `lam`) and
  * one all-reduce (sum + count) for the ensemble mean of the Lyapunov spectrum.
Both go through `torch.distributed` (NCCL over NVLink on the GPU box, gloo in the CPU tests).

Partition: round-robin blocks of `block` trajectories.  Benchmark ensembles are cost-ordered (step count grows with
the eccentricity), so contiguous halves would leave one GPU with all the expensive trajectories; interleaved blocks
give every rank the same step-count distribution while keeping neighbouring lanes of a warp similar.
"""
from __future__ import annotations

import os
from typing import Callable, Optional, Tuple

import numpy as np

from .api import EnsembleSolution


def auto_block(N: int, world: int, block: int = 4096) -> int:
    """Block size that gives every rank at least ~8 blocks: `block` for large ensembles, smaller for small ones (a 2^14
    TaylorN ensemble over 8 ranks with blocks of 4096 would leave four ranks empty)."""
    return int(max(1, min(block, -(-N // (8 * world)))))


def shard_indices(N: int, world: int, rank: int, block: int = 4096) -> np.ndarray:
    """Global trajectory indices owned by `rank`: blocks rank, rank+world, rank+2*world, ... of `block` trajectories."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    if block <= 0:
        raise ValueError("block must be positive")
    nblk = (N + block - 1) // block
    mine = np.arange(rank, nblk, world, dtype=np.int64)
    idx = (mine[:, None] * block + np.arange(block, dtype=np.int64)[None, :]).reshape(-1)
    return idx[idx < N]


def shard_sizes(N: int, world: int, block: int = 4096) -> np.ndarray:
    return np.array([shard_indices(N, world, r, block).size for r in range(world)], dtype=np.int64)


def _dist():
    import torch.distributed as dist
    if not dist.is_available() or not dist.is_initialized():
        raise RuntimeError("torch.distributed is not initialised (launch with torchrun, one process per GPU)")
    return dist


def _gather_rows(local: np.ndarray, sizes: np.ndarray, device, group) -> list:
    """All-gather of a [rows, n_local] array whose last dimension differs per rank (padded to the largest shard)."""
    import torch
    dist = _dist()
    world = sizes.size
    nmax = int(sizes.max()) if world else 0
    rows = local.shape[0]
    buf = torch.zeros((rows, nmax), dtype=torch.from_numpy(local[:0]).dtype, device=device)
    buf[:, : local.shape[1]] = torch.from_numpy(np.ascontiguousarray(local)).to(device)
    out = torch.empty((world * rows, nmax), dtype=buf.dtype, device=device)  # concatenated along dim 0
    dist.all_gather_into_tensor(out, buf, group=group)
    out = out.cpu().numpy().reshape(world, rows, nmax)
    return [out[r, :, : int(sizes[r])] for r in range(world)]


def gather_ensemble(local: EnsembleSolution, N: int, block: int = 4096, device=None, group=None) -> EnsembleSolution:
    """Reassembles the global `EnsembleSolution` (final-state fields) from every rank's shard, on every rank."""
    dist = _dist()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(N, world, block)
    if local.nsteps.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.nsteps.shape[0]} trajectories, its shard has {sizes[rank]}")
    if device is None:
        import torch
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else "cpu"
    dof = local.xfinal.shape[0]
    f64 = [local.xfinal, local.tfinal[None, :]]
    if local.lam is not None:
        f64.append(local.lam)
    packed = np.concatenate(f64, axis=0)
    ints = np.stack([local.nsteps.astype(np.int64), local.status.astype(np.int64)])
    g_f = _gather_rows(packed, sizes, device, group)
    g_i = _gather_rows(ints, sizes, device, group)
    xfinal, tfinal = np.empty((dof, N)), np.empty(N)
    nsteps, status = np.empty(N, dtype=np.int64), np.empty(N, dtype=np.int32)
    lam = np.empty((local.lam.shape[0], N)) if local.lam is not None else None
    for r in range(world):
        idx = shard_indices(N, world, r, block)
        xfinal[:, idx], tfinal[idx] = g_f[r][:dof], g_f[r][dof]
        if lam is not None:
            lam[:, idx] = g_f[r][dof + 1:]
        nsteps[idx], status[idx] = g_i[r][0], g_i[r][1].astype(np.int32)
    return EnsembleSolution(xfinal, tfinal, nsteps, status, lam=lam)


def lyap_ensemble_mean(lam_local: np.ndarray, device=None, group=None) -> np.ndarray:
    """Mean Lyapunov spectrum over the whole ensemble: one all-reduce of (sum over the shard, count)."""
    import torch
    dist = _dist()
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else "cpu"
    acc = torch.from_numpy(np.concatenate([lam_local.sum(axis=1), [float(lam_local.shape[1])]])).to(device)
    dist.all_reduce(acc, op=dist.ReduceOp.SUM, group=group)
    acc = acc.cpu().numpy()
    return acc[:-1] / acc[-1]


def run_sharded(integrate_local: Callable[[np.ndarray, Optional[np.ndarray]], EnsembleSolution], q0: np.ndarray,
                params: Optional[np.ndarray] = None, block: int = 4096, device=None, group=None) -> EnsembleSolution:
    """Every rank holds the same global `q0[dof, N]` (and `params[npar]` or `[npar, N]`); each integrates its shard
    with `integrate_local(q0_shard, params_shard)` and the results are all-gathered into global order."""
    dist = _dist()
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    q0 = np.asarray(q0, dtype=np.float64)
    N = q0.shape[1]
    block = auto_block(N, world, block)  # small ensembles: smaller blocks, so that no rank stays empty
    idx = shard_indices(N, world, rank, block)
    p = None if params is None else np.asarray(params, dtype=np.float64)
    p_loc = p if p is None or p.ndim == 1 else np.ascontiguousarray(p[:, idx])
    local = integrate_local(np.ascontiguousarray(q0[:, idx]), p_loc)
    return gather_ensemble(local, N, block, device, group)


def _local_rank_device() -> Tuple[int, int, int]:
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def taylorinteg_sharded(f, q0, t0, tmax, order, abstol, params=None, block: int = 4096, **kw) -> EnsembleSolution:
    """`taylorinteg` over an ensemble sharded across the GPUs of one node (one process per GPU, torchrun)."""
    from . import _lib as L
    from .api import taylorinteg
    L.check(L.lib().tb_set_device(_local_rank_device()[2]), "tb_set_device")
    return run_sharded(lambda q, p: taylorinteg(f, q, t0, tmax, order, abstol, p, output="final", **kw), q0, params, block)


def lyap_taylorinteg_sharded(f, q0, t0, tmax, order, abstol, params=None, block: int = 4096, **kw):
    """Sharded `lyap_taylorinteg`; returns (EnsembleSolution with `lam[dof, N]`, ensemble-mean spectrum [dof])."""
    from . import _lib as L
    from .api import lyap_taylorinteg
    L.check(L.lib().tb_set_device(_local_rank_device()[2]), "tb_set_device")
    sol = run_sharded(lambda q, p: lyap_taylorinteg(f, q, t0, tmax, order, abstol, p, output="final", **kw), q0, params,
                      block)
    return sol, sol.lam.mean(axis=1)
