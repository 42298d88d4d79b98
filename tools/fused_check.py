"""torchrun helper: fused multi-GPU peel against the oracle (N small) and against the NCCL path."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from evox_b200.distributed import FusedShardedNonDominatedSort, ShardedNonDominatedSort
from oracle import oracle_c as C
ok = True
for n, m in [(20000, 3), (5000, 10), (1024, 2), (40000, 3)]:
    x = np.random.default_rng(n + m).random((n, m), dtype=np.float32)
    x[:100] = x[100:200]
    xd = torch.from_numpy(x).cuda()
    f = FusedShardedNonDominatedSort(n, m)
    for rep in range(2):
        got = f(xd).cpu().numpy()
    ref = C.non_dominated_sort(x) if rank == 0 else None
    nc = ShardedNonDominatedSort(n, m)(xd).cpu().numpy()
    same = (got == nc).all()
    if rank == 0:
        good = (got == ref).all() and f.num_fronts == int(ref.max()) + 1
        ok &= bool(good)
        print(f"n={n} m={m} world={world} fused==oracle {good} fused==nccl {same} fronts {f.num_fronts}", flush=True)
    ok &= bool(same)
    del f
print(rank, "ALL OK" if ok else "MISMATCH", flush=True)
dist.destroy_process_group()
