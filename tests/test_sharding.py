"""Multi-GPU host logic on CPU: world_size-2 gloo process group.

The per-shard integrator is injected; here it is the CPU oracle (tests may call it), on the GPU box it is the fused kernel
(test_two_rank_sharded_taylorinteg_on_gpu below covers that through the same code path).
"""
import os
import socket

import numpy as np
import pytest

import taylorintegration_b200 as tb
from taylorintegration_b200 import problems as P, sharding as S


def test_shard_indices_partition_and_balance():
    for N, world, block in [(10, 2, 4), (4096 * 5 + 17, 4, 4096), (7, 8, 1), (0, 2, 4), (100, 1, 16)]:
        parts = [S.shard_indices(N, world, r, block) for r in range(world)]
        allidx = np.concatenate(parts) if parts else np.empty(0)
        assert np.array_equal(np.sort(allidx), np.arange(N))
        assert np.array_equal(S.shard_sizes(N, world, block), [p.size for p in parts])
        assert max(p.size for p in parts) - min(p.size for p in parts) <= block
    with pytest.raises(ValueError):
        S.shard_indices(10, 2, 2)
    with pytest.raises(ValueError):
        S.shard_indices(10, 2, 0, block=0)


def test_interleaved_blocks_balance_cost_ordered_ensemble():
    # step counts grow with e: interleaved blocks must give both ranks (nearly) the same work
    N = 1 << 16
    cost = 320 + 5000 * (np.arange(N) / (N - 1)) ** 3
    w = [cost[S.shard_indices(N, 2, r, 1024)].sum() for r in range(2)]
    assert abs(w[0] - w[1]) / sum(w) < 0.03
    contiguous = [cost[: N // 2].sum(), cost[N // 2:].sum()]
    assert abs(contiguous[0] - contiguous[1]) / sum(contiguous) > 0.3


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, outdir: str):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import oracle
    import taylorintegration_b200 as tb  # noqa: F811
    from taylorintegration_b200 import _lib as L, problems as P, sharding as S  # noqa: F811
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # plain driver, shared parameter
        tape = P.kepler_tape()
        q0, mu, _ = P.kepler_ensemble(37)
        tmax = 2 * np.pi
        sol = S.run_sharded(lambda q, p: oracle.integrate(tape, 28, q, 0.0, tmax, p, maxsteps=2000), q0, mu, block=4)
        # Lyapunov driver, swept parameter + ensemble mean through the all-reduce
        lt = P.lorenz_tape()
        N = 5
        par = np.stack([np.full(N, 16.0), 40.0 + np.arange(N), np.full(N, 4.0)])
        lq = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
        lsol = S.run_sharded(lambda q, p: oracle.integrate(lt, 20, q, 0.0, 0.5, p, kind=L.KIND_LYAP, maxsteps=2000), lq, par,
                             block=2)
        idx = S.shard_indices(N, world, rank, 2)
        mean = S.lyap_ensemble_mean(lsol.lam[:, idx])
        np.savez(os.path.join(outdir, f"r{rank}.npz"), x=sol.xfinal, t=sol.tfinal, n=sol.nsteps, s=sol.status, lam=lsol.lam,
                 lmean=mean, ln=lsol.nsteps)
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_gather_matches_single_process(tmp_path):
    import torch.multiprocessing as mp
    import oracle
    from taylorintegration_b200 import _lib as L
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    tape = P.kepler_tape()
    q0, mu, _ = P.kepler_ensemble(37)
    ref = oracle.integrate(tape, 28, q0, 0.0, 2 * np.pi, mu, maxsteps=2000)
    lt = P.lorenz_tape()
    N = 5
    par = np.stack([np.full(N, 16.0), 40.0 + np.arange(N), np.full(N, 4.0)])
    lq = np.tile(np.array([[19.0], [20.0], [50.0]]), (1, N))
    lref = oracle.integrate(lt, 20, lq, 0.0, 0.5, par, kind=L.KIND_LYAP, maxsteps=2000)
    for r in range(2):
        z = np.load(tmp_path / f"r{r}.npz")
        assert np.array_equal(z["x"], ref.xfinal) and np.array_equal(z["t"], ref.tfinal)
        assert np.array_equal(z["n"], ref.nsteps) and np.array_equal(z["s"], ref.status)
        assert np.array_equal(z["lam"], lref.lam) and np.array_equal(z["ln"], lref.nsteps)
        assert np.allclose(z["lmean"], lref.lam.mean(axis=1), rtol=1e-14, atol=0)


def test_gather_requires_process_group():
    sol = tb.EnsembleSolution(np.zeros((1, 1)), np.zeros(1), np.zeros(1, dtype=np.int64), np.zeros(1, dtype=np.int32))
    with pytest.raises(RuntimeError):
        S.gather_ensemble(sol, 1)


def _gpu_worker(rank: int, world: int, port: int, outdir: str):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch.distributed as dist
    from taylorintegration_b200 import problems as P, sharding as S  # noqa: F811
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK="0")
    dist.init_process_group("gloo", rank=rank, world_size=world)  # both ranks share cuda:0; gloo carries the gather
    try:
        q0, mu, _ = P.kepler_ensemble(300)
        sol = S.taylorinteg_sharded(P.kepler_tape(), q0, 0.0, 2 * np.pi, 28, 1e-20, mu, block=32, maxsteps=2000)
        np.savez(os.path.join(outdir, f"g{rank}.npz"), x=sol.xfinal, n=sol.nsteps, s=sol.status)
    finally:
        dist.destroy_process_group()


@pytest.mark.gpu
def test_two_rank_sharded_taylorinteg_on_gpu(tmp_path):
    import torch.multiprocessing as mp
    import oracle
    mp.spawn(_gpu_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    q0, mu, _ = P.kepler_ensemble(300)
    ref = oracle.integrate(P.kepler_tape(), 28, q0, 0.0, 2 * np.pi, mu, maxsteps=2000)
    for r in range(2):
        z = np.load(tmp_path / f"g{r}.npz")
        assert np.array_equal(z["x"], ref.xfinal) and np.array_equal(z["n"], ref.nsteps) and not z["s"].any()


def test_auto_block_leaves_no_rank_empty():
    # 2^14 TaylorN neighbourhoods over 8 ranks: blocks of 4096 would feed four ranks only
    assert S.auto_block(1 << 20, 8) == 4096 and S.auto_block(1 << 14, 8) == 256 and S.auto_block(5, 8) == 1
    for N, world in ((1 << 14, 8), (2048, 8), (100, 3), (8, 8)):
        sizes = S.shard_sizes(N, world, S.auto_block(N, world))
        assert sizes.sum() == N and sizes.min() >= 1 and sizes.max() - sizes.min() <= S.auto_block(N, world)
