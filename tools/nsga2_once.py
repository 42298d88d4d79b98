"""A few NSGA-II generations at pop 100k on DTLZ2 (for launch lists / stage timings)."""
import sys; sys.path.insert(0, '.')
import torch, evox_b200 as E
from evox_b200.algorithms import NSGA2, run_generations
alg = NSGA2(torch.zeros(12), torch.ones(12), n_objs=3, pop_size=100_000)
prob = E.problems.numerical.DTLZ2(d=12, m=3)
state, _ = run_generations(alg, prob, alg.setup(seed=0), 30)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); state, _ = run_generations(alg, prob, state, 20); e1.record(); torch.cuda.synchronize()
print("ms per generation", e0.elapsed_time(e1) / 20, "fronts", int(E.lib().evx_nds_last_num_fronts()))
