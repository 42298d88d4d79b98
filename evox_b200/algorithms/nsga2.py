"""NSGA-II generation loop over the CUDA hot path (mirror of src/evox/algorithms/mo/nsga2.py:22-96
driven the way StdWorkflow.step does, src/evox/workflows/std_workflow.py:187-200).

In production the reference's own NSGA2 / StdWorkflow stay the callers (they are JAX host code and
only see the unchanged `evox.operators` API).  JAX is not installable in this image, so this module
exists to (i) measure NSGA-II generations/s, the second half of BASELINE.json's metric, and (ii)
serve users without JAX.  `ask` is one CUDA kernel (SBX + polynomial mutation + clip); `tell` is
concat + NonDominate exactly as nsga2.py:89-96.
"""
import torch

from .._dev import as_f32_cuda, stream_ptr
from .._lib import check, lib
from ..core import State
from ..operators.selection import NonDominate


class NSGA2:
    def __init__(self, lb, ub, n_objs, pop_size, pro_c=1.0, dis_c=20.0, pro_m=1.0, dis_m=20.0):
        self.lb = as_f32_cuda(lb, "lb")
        self.ub = as_f32_cuda(ub, "ub")
        self.n_objs, self.pop_size, self.dim = n_objs, pop_size, self.lb.shape[0]
        self.pro_c, self.dis_c, self.pro_m, self.dis_m = pro_c, dis_c, pro_m, dis_m     # sbx.py:55, pm_mutation.py:67
        self.survivor_selection = NonDominate(pop_size)

    def setup(self, seed=0):
        g = torch.Generator(device="cuda").manual_seed(int(seed))
        pop = torch.rand((self.pop_size, self.dim), device="cuda", generator=g) * (self.ub - self.lb) + self.lb
        return State(population=pop, fitness=torch.zeros((self.pop_size, self.n_objs), device="cuda"),
                     next_generation=pop, seed=int(seed), generation=0)

    def init_ask(self, state):
        return state.population, state

    def init_tell(self, state, fitness):
        return state.replace(fitness=fitness)

    def ask(self, state):
        pop = state.population
        out = torch.empty_like(pop)
        check(lib().evx_variation_sbx_pm_f32(pop.data_ptr(), pop.shape[0], pop.shape[1], self.lb.data_ptr(),
                                             self.ub.data_ptr(), self.pro_c, self.dis_c, self.pro_m, self.dis_m,
                                             state.seed & 0xFFFFFFFF, state.generation & 0xFFFFFFFF, out.data_ptr(),
                                             stream_ptr()))
        return out, state.replace(next_generation=out, generation=state.generation + 1)

    def tell(self, state, fitness):
        merged_pop = torch.cat([state.population, state.next_generation], dim=0)
        merged_fit = torch.cat([state.fitness, fitness], dim=0)
        survivor, survivor_fit = self.survivor_selection(merged_pop, merged_fit)
        return state.replace(population=survivor, fitness=survivor_fit)


def run_generations(algorithm, problem, state, generations, pstate=None):
    """init_ask/init_tell once (std_workflow.py:204-214) if the state is fresh, then ask-evaluate-tell."""
    if state.generation == 0:
        pop, state = algorithm.init_ask(state)
        fit, pstate = problem.evaluate(pstate, pop)
        state = algorithm.init_tell(state, fit)
    for _ in range(generations):
        cand, state = algorithm.ask(state)
        fit, pstate = problem.evaluate(pstate, cand)
        state = algorithm.tell(state, fit)
    return state, pstate
