"""Small calls of the single-CTA kernels (k_small_sort_unique, k_peel_single, k_niche_loop, k_spea2_*), for
`compute-sanitizer --tool racecheck python tools/race_case.py` (shared-memory hazards) and `--tool memcheck`."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import evox_b200 as E
from evox_b200.algorithms import nsga3, spea2
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.rand(2600, 3, device="cuda", generator=g); x[:100] = x[100:200]
r = E.operators.non_dominated_sort(x)                                        # 3 panels: W = 16, RS = 2
y = torch.rand(5000, 3, device="cuda", generator=g)
r5 = E.operators.non_dominated_sort(y)                                       # 5 panels: W = 8, RS = 1
z = torch.rand(2500, 2, device="cuda", generator=g); z[:, 1] = 1 - z[:, 0]   # one front: two gather passes
rz = E.operators.non_dominated_sort(z)
n, R, ps = 3000, 7, 1500
rank = torch.sort(torch.randint(0, 6, (n,), device="cuda", generator=g)).values.int()[torch.randperm(n, device="cuda", generator=g)]
last = int(rank[torch.sort(rank, stable=True).indices[ps]])
gid = torch.randint(0, R, (n,), device="cuda", generator=g).int(); gid[rank > last] = R
gd = (torch.randint(0, 50, (n,), device="cuda", generator=g) / 50).float(); gd[rank > last] = float("nan")
rw = torch.randint(0, 2**31 - 1, (ps + 1, 2), device="cuda", dtype=torch.int32, generator=g)
sel = nsga3.niche(rank, last, gid, gd, R, ps, rw)                           # heavy niches: one warp each
sel2 = nsga3.niche(rank, last, torch.randint(0, 900, (n,), device="cuda", generator=g).int(), gd.nan_to_num(0.3), 900, ps, rw)
obj = torch.rand(1500, 3, device="cuda", generator=g)
fit = spea2.cal_fitness(obj)
keep = spea2._truncation(obj, 600, torch.ones(1500, dtype=torch.bool, device="cuda"))
torch.cuda.synchronize()
print("race_case ok", int(r.max()) + 1, int(r5.max()) + 1, int(rz.max()) + 1, int(sel.numel()), int(keep.sum()))
