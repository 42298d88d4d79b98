"""Device times of the NSGA-III selection path at the configs[3] shape (merged N = 100k, M = 10, 48 620 reference
directions = Das-Dennis H = 9, pop 50k): association, niching, and the whole environmental selection."""
import sys, numpy as np, torch
sys.path.insert(0, '.')
from evox_b200.algorithms import nsga3
from evox_b200.operators.selection.non_dominate import non_dominated_sort


def med(fn, reps=5):
    for _ in range(2):
        fn()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


for n, m, R, ps in [(100_000, 10, 48_620, 50_000), (20_000, 3, 10_011, 10_000)]:
    g = torch.Generator(device="cuda").manual_seed(3)
    fit = torch.rand((n, m), device="cuda", generator=g)
    ref = torch.rand((R, m), device="cuda", generator=g) + 0.01
    rank = non_dominated_sort(fit)
    order = torch.sort(rank, stable=True).indices
    last_rank = int(rank[order[ps]])
    norm = nsga3.normalize(fit, rank, last_rank)
    gid, gd = nsga3.associate(norm, ref)
    rw = torch.randint(0, 2**31 - 1, (ps + 1, 2), device="cuda", dtype=torch.int32, generator=g)
    ta = med(lambda: nsga3.associate(norm, ref))
    tn = med(lambda: nsga3.niche(rank, last_rank, gid, gd, R, ps, rw))
    pop = torch.rand((n, 19), device="cuda", generator=g)
    te = med(lambda: nsga3.environmental_selection(pop, fit, ref, ps, rw))
    picks = ps - int((rank < last_rank).sum())
    print(f"n {n} m {m} R {R} pop {ps}: last_rank {last_rank} picks in the niching loop {picks}; associate {ta:.3f} ms "
          f"({n * R / ta / 1e6:.1f} G pairs/s), niche {tn:.3f} ms, whole selection {te:.3f} ms")
