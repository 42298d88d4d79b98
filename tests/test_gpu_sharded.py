"""CUDA side of the sharded non-dominated sort (evx_nds_shard_*), on ONE GPU: the ranks of a
world of 1, 2, 3 and 8 are driven one after another from a single process (plain sequential
launches, nothing waits on another kernel), their masks are concatenated where NCCL would
all-gather them, and the summed partial ranks must equal the oracle's."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run_world(x, world):
    from evox_b200.distributed import _CudaShard
    n, m = x.shape
    shards = [_CudaShard(n, m, world, r) for r in range(world)]
    mw = shards[0].mask_words
    assert all(s.mask_words == mw for s in shards)
    local = [torch.zeros(mw, dtype=torch.int32, device="cuda") for _ in range(world)]
    anyf = torch.zeros(1, dtype=torch.int32, device="cuda")
    for s in shards:
        s.build(x)
    k = 0
    while True:
        for s, lm in zip(shards, local):
            s.detect(k, lm)
        allm = torch.cat(local)                      # what all_gather_into_tensor produces
        for s in shards:
            s.apply(k, allm, anyf)
        if int(anyf.item()) == 0:
            break
        k += 1
        assert k < n + 2
    total = torch.zeros(n, dtype=torch.int32, device="cuda")
    part = torch.zeros(n, dtype=torch.int32, device="cuda")
    for s in shards:
        s.ranks(part)
        total += part                                # what all_reduce(SUM) produces
    return total, k


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("n,m", [(9000, 3), (2500, 10)])
def test_shard_kernels_emulated_world(world, n, m):
    from oracle import oracle_c as C
    rng = np.random.default_rng(world * 100 + m)
    x = rng.random((n, m), dtype=np.float32)
    x[:300] = x[300:600]
    x[11, 1] = np.nan
    ref = C.non_dominated_sort(x)
    got, fronts = _run_world(torch.from_numpy(x).cuda(), world)
    assert (got.cpu().numpy() == ref).all()
    assert fronts == int(ref.max()) + 1


def test_sharded_driver_single_process():
    """ShardedNonDominatedSort without an initialised process group = world of one (stepwise peel)."""
    from oracle import oracle_c as C
    from evox_b200.distributed import ShardedNonDominatedSort
    x = np.random.default_rng(1).random((12000, 3), dtype=np.float32)
    sh = ShardedNonDominatedSort(12000, 3)
    ref = C.non_dominated_sort(x)
    for _ in range(2):                               # plan and workspace are reusable
        got = sh(torch.from_numpy(x).cuda())
        assert (got.cpu().numpy() == ref).all() and sh.num_fronts == int(ref.max()) + 1


def test_stepwise_peel_env_switch(monkeypatch):
    """EVX_PEEL=stepwise selects the per-front launches inside evx_nds_f32 (the persistent kernel is
    the default); both give the same ranks."""
    import evox_b200 as E
    x = torch.rand(30000, 3, device="cuda")
    a = E.operators.non_dominated_sort(x)
    monkeypatch.setenv("EVX_PEEL", "stepwise")
    b = E.operators.non_dominated_sort(x)
    assert torch.equal(a, b)


def test_fused_peel_world_of_one():
    """The fused multi-rank peel kernel (front replica in peer-mapped memory, committers push to every
    rank, world-wide barrier) with a process group of one rank: same kernel and protocol as on 8 GPUs,
    peers == self.  Three calls in a row on one plan: the peer state is cleared once and the barrier
    epochs run on across launches."""
    import os
    import torch.distributed as dist
    from oracle import oracle_c as C
    from evox_b200.distributed import FusedShardedNonDominatedSort
    created = False
    if not dist.is_initialized():
        port = 29600 + os.getpid() % 300
        dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=0, world_size=1,
                                device_id=torch.device("cuda", torch.cuda.current_device()))
        created = True
    try:
        try:
            f = FusedShardedNonDominatedSort(30000, 3)
        except (ImportError, RuntimeError, AttributeError) as e:     # no symmetric memory on this box
            pytest.skip(f"torch symmetric memory unavailable: {e}")
        rng = np.random.default_rng(6)
        for rep in range(3):
            x = rng.random((30000, 3), dtype=np.float32)
            x[:500] = x[500:1000]
            ref = C.non_dominated_sort(x)
            got = f(torch.from_numpy(x).cuda())
            assert (got.cpu().numpy() == ref).all() and f.num_fronts == int(ref.max()) + 1, rep
    finally:
        if created:
            dist.destroy_process_group()


def _free_port():
    import socket
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_peel_real_ranks(world):
    """One process per GPU under torch.distributed.run (skipped on a box with fewer GPUs): the fused peel
    against the oracle, against the NCCL-exchange path, and - at n = 200 000 - against the sampled
    exactness check, several calls per plan (tools/fused_check.py holds the comparison)."""
    import subprocess
    import sys
    from pathlib import Path
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs, this box has {torch.cuda.device_count()}")
    root = Path(__file__).resolve().parent.parent
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(root / "tools" / "fused_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ALL OK") == world and "MISMATCH" not in r.stdout, r.stdout[-3000:]
