"""The N>1 path of the sharded non-dominated sort on CPU: two gloo ranks run the real driver
(evox_b200/distributed.py) over a NumPy stand-in for the device side (tests/_np_shard.py), and the
assembled ranks must equal the oracle's bit for bit."""
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _worker(rank, world, port, n, m, seed, q):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from _np_shard import NumpyShard
        from evox_b200.distributed import ShardedNonDominatedSort
        rng = np.random.default_rng(seed)
        x = rng.random((n, m), dtype=np.float32)
        x[: n // 10] = x[n // 10: 2 * (n // 10)]          # duplicates
        x[3, 0] = np.nan
        sh = ShardedNonDominatedSort(n, m, shard_factory=NumpyShard)
        out = sh(torch.from_numpy(x)).numpy().copy()
        q.put((rank, out, sh.num_fronts, sh.collectives))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,m", [(2, 2600, 3), (2, 1500, 2), (3, 3300, 3)])
def test_sharded_driver_matches_oracle(world, n, m):
    from oracle import oracle_np as O
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() + n) % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, m, 5, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    rng = np.random.default_rng(5)
    x = rng.random((n, m), dtype=np.float32)
    x[: n // 10] = x[n // 10: 2 * (n // 10)]
    x[3, 0] = np.nan
    ref = O.non_dominated_sort(x)
    for rank, out, fronts, coll in results:
        assert (out == ref).all(), rank                       # every rank ends with the full ranks
        assert fronts == int(ref.max()) + 1
        assert coll >= fronts + 1                            # one all-gather per front + final all-reduce


def test_panel_ownership_covers_every_column_once():
    sys.path.insert(0, str(ROOT / "tests"))
    from _np_shard import NumpyShard, TW
    for world in (1, 2, 4, 8):
        n = 10_000
        owners = np.zeros(n, int)
        for r in range(world):
            s = NumpyShard(n, 3, world, r)
            owners += s._owned(np.arange(n))
            assert s.mask_words * 32 >= int(s._owned(np.arange(n)).sum())
        assert (owners == 1).all()


# ---- the evaluation shard of the N-GPU NSGA-II generation (std_workflow.py:187-200 + core/distributed.py:49-56) ----
class _CpuProblem:
    """Stand-in problem on CPU tensors (the CUDA evaluate is covered by the -m gpu tests): a row-wise
    function, so slicing rows and gathering must reproduce the unsharded result exactly."""

    def evaluate(self, state, X):
        return torch.stack([X.sum(1), (X * X).sum(1), X[:, 0] - X[:, -1]], dim=1), state


def _eval_worker(rank, world, port, n, d, q):
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from evox_b200.distributed import recommended_world, sharded_evaluate
        X = torch.from_numpy(np.random.default_rng(9).random((n, d), dtype=np.float32))
        state = object()
        F, st = sharded_evaluate(_CpuProblem(), state, X)
        q.put((rank, F.numpy().copy(), st is state, recommended_world(200_000), recommended_world(1_000_000, world)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 1000), (2, 1001), (3, 10), (2, 1)])
def test_sharded_evaluate_matches_unsharded(world, n):
    """Rows dealt in equal slices, all-gather of the fitness; a population that does not divide by the world
    size (ragged last slice, even an empty one) gives the same F on every rank as one evaluate call."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + (os.getpid() + n + world) % 2000
    procs = [ctx.Process(target=_eval_worker, args=(r, world, port, n, 5, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    X = torch.from_numpy(np.random.default_rng(9).random((n, 5), dtype=np.float32))
    want, _ = _CpuProblem().evaluate(None, X)
    for rank, F, same_state, w_small, w_big in results:
        assert F.shape == (n, 3) and (F == want.numpy()).all(), rank
        assert same_state and w_small == 1 and w_big == (2 if world >= 2 else 1)
