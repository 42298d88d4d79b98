import sys, torch, numpy as np
sys.path.insert(0, '/root/repo')
import evox_b200 as E
from evox_b200._lib import lib, check
L = lib()
def t_of(fn, reps=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3
for n in (200, 1000, 2048, 4096):
    x = torch.rand(n, 3, device="cuda")
    r = torch.empty(n, dtype=torch.int32, device="cuda")
    a = t_of(lambda: E.operators.non_dominated_sort(x))
    b = t_of(lambda: check(L.evx_nds_noalloc_f32(x.data_ptr(), n, 3, r.data_ptr(), None)))
    print(f"n={n:5d}  workspace path {a:8.1f} us   small (3 launches, no sync) {b:8.1f} us")
