"""Device times of SPEA2's fitness assignment and archive truncation (merged population n, M = 3)."""
import sys, torch
sys.path.insert(0, '.')
from evox_b200.algorithms import spea2


def med(fn, reps=3):
    fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


for n, m in [(20_000, 3), (100_000, 3), (20_000, 10)]:
    g = torch.Generator(device="cuda").manual_seed(3)
    obj = torch.rand((n, m), device="cuda", generator=g)
    tf = med(lambda: spea2.cal_fitness(obj))
    mask = torch.rand(n, device="cuda", generator=g) < 0.6                 # 60 % valid, cut to 50 %
    k = int(mask.sum()) - n // 2
    tt = med(lambda: spea2._truncation(obj, k, mask))
    print(f"n {n} m {m}: cal_fitness {tf:.3f} ms ({n * n / tf / 1e6:.1f} G pairs/s); truncation of {k} rows {tt:.3f} ms")
