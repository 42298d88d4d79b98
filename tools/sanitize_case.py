"""One small call of every entry point, for `compute-sanitizer --tool memcheck python tools/sanitize_case.py`."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import evox_b200 as E
from evox_b200._lib import lib, check
torch.manual_seed(0)
x = torch.rand(3000, 3, device="cuda"); x[:100] = x[100:200]
r = E.operators.non_dominated_sort(x)
r2 = E.operators.non_dominated_sort(torch.rand(700, 3, device="cuda"))          # small path
r10 = E.operators.non_dominated_sort(torch.rand(2100, 10, device="cuda"))
r13 = E.operators.non_dominated_sort(torch.rand(1100, 13, device="cuda"))         # generic tile kernel
cd = E.operators.crowding_distance(x, r == 2)
cs = E.operators.crowding_distance_sort(x)
d = E.utils.dominate_relation(x[:500], x[:700]); db = E.utils.dominate_relation_bits(x[:500], x[:700])
from evox_b200.operators.selection.non_dominate import host_rank_from_domination_matrix
hr = host_rank_from_domination_matrix(E.utils.dominate_relation(x[:600], x[:600]), None)
p, f = E.operators.selection.non_dominate(torch.rand(3000, 5, device="cuda"), x, 1500, False)
for k in range(1, 8):
    getattr(E.problems.numerical, f"DTLZ{k}")(d=12, m=3).evaluate(None, torch.rand(1000, 12, device="cuda"))
for k in range(1, 10):
    getattr(E.problems.numerical, f"LSMOP{k}")(d=300, m=3).evaluate(None, torch.rand(200, 300, device="cuda"))
E.problems.numerical.MaF1(d=19, m=10).evaluate(None, torch.rand(500, 19, device="cuda"))
from evox_b200.algorithms import NSGA2, run_generations
alg = NSGA2(torch.zeros(12), torch.ones(12), 3, 1500)
st, _ = run_generations(alg, E.problems.numerical.DTLZ2(d=12, m=3), alg.setup(1), 2)
rk = torch.empty(900, dtype=torch.int32, device="cuda")
check(lib().evx_nds_noalloc_f32(x.data_ptr(), 900, 3, rk.data_ptr(), None))
from evox_b200.distributed import ShardedNonDominatedSort
s1 = ShardedNonDominatedSort(3000, 3)(x)
torch.cuda.synchronize()
assert torch.equal(s1, r)
print("sanitize_case ok", int(r.max()) + 1, int(r10.max()) + 1)
