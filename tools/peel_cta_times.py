"""Distribution over CTAs of the peel's phase clocks (one GPU, shard plan with world = 1 and the fused kernel's
single-rank twin): who is the slowest CTA of a front, and by how much."""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, '.')
from evox_b200._lib import lib, check
from evox_b200.distributed import _CudaShard
L = lib()
for n in [int(a) for a in sys.argv[1:]] or [200_000]:
    x = torch.from_numpy(np.random.default_rng(0).random((n, 3), dtype=np.float32)).cuda()
    sh = _CudaShard(n, 3, 1, 0)
    ptrs = (ctypes.c_void_p * 1)()
    peer = torch.zeros(int(L.evx_nds_shard_peer_bytes(sh.plan)), dtype=torch.uint8, device="cuda")
    ptrs[0] = peer.data_ptr()
    s = torch.cuda.current_stream().cuda_stream
    check(L.evx_nds_shard_peer_init(sh.plan, peer.data_ptr(), s))
    for _ in range(3):
        sh.build(x)
        check(L.evx_nds_shard_peel_fused(sh.plan, ptrs, None, s))
    torch.cuda.synchronize()
    fronts = int(L.evx_nds_last_num_fronts())
    buf = (ctypes.c_uint64 * (4 * 1024))()
    check(L.evx_debug_peel_cta_times(sh.plan, buf, 1024))
    a = np.array(buf[:], dtype=np.float64).reshape(1024, 4) / fronts / 1e3
    a = a[a.sum(1) > 0]
    busy = a[:, 0] + a[:, 1] + a[:, 3]
    q = lambda v: "min %.2f  p50 %.2f  p90 %.2f  max %.2f" % (v.min(), np.median(v), np.quantile(v, 0.9), v.max())
    print(f"n {n} fronts {fronts} CTAs {len(a)} (us per front)")
    for nm, v in (("gather", a[:, 0]), ("merge+commit+publish", a[:, 1]), ("scans", a[:, 3]), ("busy = sum of those", busy), ("waiting", a[:, 2])):
        print(f"  {nm:22s} {q(v)}")
    print(f"  period {np.mean(busy + a[:, 2]):.2f}; slowest CTA's busy time covers {busy.max() / np.mean(busy + a[:, 2]) * 100:.0f} % of it")
