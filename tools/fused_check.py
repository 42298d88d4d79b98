"""torchrun helper (also run by tests/test_gpu_sharded.py::test_fused_peel_real_ranks): the fused multi-GPU
peel against the oracle (small n), against the NCCL-exchange path, and against the sampled exactness check
(n = 200 000), several calls per plan.  Every rank prints "<rank> ALL OK" or "<rank> MISMATCH"; exit code 1
on any mismatch."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from evox_b200._verify import fronts_contiguous, rank_mismatches_sampled  # noqa: E402
from evox_b200.distributed import FusedShardedNonDominatedSort, ShardedNonDominatedSort  # noqa: E402
from oracle import oracle_c as C  # noqa: E402

ok = True
QUICK = "--quick" in sys.argv                              # only the small cases (ranks without panels included)
for n, m in [(5000, 10), (3000, 3)] if QUICK else [(20000, 3), (5000, 10), (1024, 2), (40000, 3)]:
    f = FusedShardedNonDominatedSort(n, m)
    nccl = ShardedNonDominatedSort(n, m)
    for rep in range(3):                                   # one plan, three populations: no reset in between
        x = np.random.default_rng(n + m + 1000 * rep).random((n, m), dtype=np.float32)
        x[:100] = x[100:200]
        xd = torch.from_numpy(x).cuda()
        got = f(xd).cpu().numpy()
        fronts = f.num_fronts
        same = bool((got == nccl(xd).cpu().numpy()).all())
        good = True
        if rank == 0:
            ref = C.non_dominated_sort(x)
            good = bool((got == ref).all()) and fronts == int(ref.max()) + 1
            print(f"n={n} m={m} world={world} rep={rep} fused==oracle {good} fused==nccl {same} fronts {fronts}", flush=True)
        ok &= good and same
    del f, nccl
if QUICK:
    flag = torch.tensor([0 if ok else 1], device="cuda")
    dist.all_reduce(flag)
    print(rank, "ALL OK" if int(flag) == 0 else "MISMATCH", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if int(flag) == 0 else 1)
n = 200_000                                                # configs[2] shape: sampled exactness, every rank its share
f = FusedShardedNonDominatedSort(n, 3)
g = torch.Generator(device="cuda").manual_seed(7)
for rep in range(2):
    x = torch.rand((n, 3), device="cuda", generator=g)
    dist.broadcast(x, 0)
    r = f(x)
    idx = torch.randperm(n, device="cuda", generator=g)[:1024]
    dist.broadcast(idx, 0)
    bad = rank_mismatches_sampled(x, r, idx[rank::world]) + (0 if fronts_contiguous(r) else 1)
    chk = torch.tensor([int(r.long().sum())], device="cuda")
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    good = bad == 0 and bool((lo == hi).all())
    if rank == 0:
        print(f"n={n} m=3 world={world} rep={rep} sampled exactness {good} fronts {f.num_fronts}", flush=True)
    ok &= good
# NSGA-II generations (configs[2] algorithm, smaller population): evaluation sharded by rows + all-gather,
# sort sharded by column panels, selection on the gathered ranks - survivors must be bit-identical to the
# single-GPU generation loop run on this rank's own GPU.
import evox_b200 as E  # noqa: E402
from evox_b200.algorithms import NSGA2, run_generations  # noqa: E402
from evox_b200.distributed import ShardedNSGA2  # noqa: E402
for pop in (20000, 10001):                                 # the second does not divide by the world size
    alg = NSGA2(torch.zeros(12), torch.ones(12), n_objs=3, pop_size=pop)
    prob = E.problems.numerical.DTLZ2(d=12, m=3)
    s1, _ = run_generations(alg, prob, alg.setup(seed=5), 4)
    sN, _ = ShardedNSGA2(alg).run(prob, alg.setup(seed=5), 4)
    good = bool(torch.equal(s1.population, sN.population)) and bool(torch.equal(s1.fitness, sN.fitness))
    if rank == 0:
        print(f"NSGA-II pop={pop} world={world}: 4 sharded generations == 4 single-GPU generations {good}", flush=True)
    ok &= good
flag = torch.tensor([0 if ok else 1], device="cuda")
dist.all_reduce(flag)
print(rank, "ALL OK" if int(flag) == 0 else "MISMATCH", flush=True)
dist.destroy_process_group()
sys.exit(0 if int(flag) == 0 else 1)
