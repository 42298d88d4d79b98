"""Phase clocks of the single-GPU peel kernel (CTA 0 and the last CTA), per front."""
import ctypes, sys, numpy as np, torch
sys.path.insert(0, '.')
from evox_b200._lib import lib, check
L = lib()
sizes = [int(a) for a in sys.argv[1:]] or [1_000_000]
m = 3
for n in sizes:
  if True:
    x = torch.from_numpy(np.random.default_rng(0).random((n, m), dtype=np.float32)).cuda()
    ws = torch.empty(L.evx_nds_workspace_bytes(n, m), dtype=torch.uint8, device="cuda")
    r = torch.empty(n, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for i in range(4):
        L.evx_profile_enable(1 if i == 3 else 0)
        check(L.evx_nds_f32(x.data_ptr(), n, m, r.data_ptr(), ws.data_ptr(), ws.numel(), s))
    a, b, c = ctypes.c_float(), ctypes.c_float(), ctypes.c_float()
    L.evx_profile_last(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
    dbg = (ctypes.c_uint64 * 16)()
    check(L.evx_debug_peel_times(None, dbg))
    print(f"n {n} prep {a.value:.3f} tiles {b.value:.3f} peel {c.value:.3f} ms")
    for nm, o in (("first CTA", 0), ("last CTA", 4)):
        k = max(1, dbg[o + 3])
        print(f"{nm}: fronts {k} per front: accumulate {dbg[o]/k/1e3:7.2f} us  merge+commit {dbg[o+1]/k/1e3:6.2f} us  wait {dbg[o+2]/k/1e3:6.2f} us (until the hint {dbg[8 + o]/k/1e3:5.2f})  scans {dbg[9 + o]/k/1e3:5.2f} us")
