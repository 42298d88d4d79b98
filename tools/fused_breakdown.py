"""torchrun helper: per-stage device times of the fused multi-GPU non-dominated sort (rank 0 prints)."""
import ctypes, os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
rank=int(os.environ["RANK"]); local=int(os.environ["LOCAL_RANK"]); world=int(os.environ["WORLD_SIZE"])
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
from evox_b200.distributed import FusedShardedNonDominatedSort
from evox_b200._lib import lib, check
from evox_b200._dev import stream_ptr
L = lib()
x = torch.from_numpy(np.random.default_rng(0).random((n, 3), dtype=np.float32)).cuda()
f = FusedShardedNonDominatedSort(n, 3)
for _ in range(3): f(x)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(6)]
acc = np.zeros(5)
R = 5
for _ in range(R):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    ev[0].record()
    f.shard.build(x)
    ev[1].record()
    ev[2].record()
    check(L.evx_nds_shard_peel_fused(f.shard.plan, f._ptrs, None, stream_ptr()))
    ev[3].record()
    f.shard.ranks(f.out)
    ev[4].record()
    dist.all_reduce(f.out)
    ev[5].record()
    torch.cuda.synchronize()
    acc += [ev[i].elapsed_time(ev[i+1]) for i in range(5)]
acc /= R
t = torch.tensor(acc, device="cuda"); mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX); mn = t.clone(); dist.all_reduce(mn, op=dist.ReduceOp.MIN)
if rank == 0:
    names = ["build(sort+tiles)", "(nothing)", "fused peel kernel", "rank scatter", "all_reduce ranks"]
    for nm, a, b in zip(names, mn.tolist(), mx.tolist()): print(f"{nm:22s} min {a:8.3f} ms  max {b:8.3f} ms")
    dbg = (ctypes.c_uint64 * 16)()
    check(L.evx_debug_peel_times(f.shard.plan, dbg))
    for nm, o in (("first CTA", 0), ("last CTA", 4)):
        k = max(1, dbg[o + 3])
        print(f"{nm}: per front  accumulate {dbg[o]/k/1e3:7.2f} us  merge+commit {dbg[o+1]/k/1e3:6.2f} us  wait for next front {dbg[o+2]/k/1e3:6.2f} us (until the hint {dbg[8 + o]/k/1e3:5.2f})  scans {dbg[9 + o]/k/1e3:5.2f} us")
    print("fronts", f.num_fronts, "world", world, "n", n, flush=True)
dist.destroy_process_group()
