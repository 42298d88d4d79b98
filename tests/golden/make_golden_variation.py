"""Fixtures for SURVEY.md section 8 row f2 (NSGA-II variation), produced by EXECUTING THE REFERENCE'S OWN
SOURCE (operators/crossover/sbx.py:8-40, operators/mutation/pm_mutation.py:7-54, read in place from
/root/reference) on the NumPy-backed jax shim, with `jax.random` replaced by a stand-in that hands out
the uniforms the CUDA kernel draws (oracle/philox.py).  Run in the build container only:

    python tests/golden/make_golden_variation.py

The reference draws from Threefry and the kernel from Philox, so no fixture can pin the random
numbers; these pin everything after the draw: for the same uniforms, the kernel must produce what the
reference's crossover -> mutation -> clip (algorithms/mo/nsga2.py:78-85) produces.
"""
import sys
from functools import partial
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parent.parent))
import jax_shim  # noqa: E402
from oracle import philox  # noqa: E402

ref = jax_shim.install()
A = ref.asarray


class FakeRandom:
    """jax.random stand-in: split() hands out (key, i) tokens, the draw functions look them up."""

    def __init__(self):
        self.table = {}

    def split(self, key, num=2):
        return [(key, i) for i in range(num)]

    def uniform(self, key, shape=(), **_kw):
        v = self.table[key]
        assert tuple(v.shape) == tuple(shape), (key, v.shape, shape)
        return A(v)

    def randint(self, key, shape=(), minval=0, maxval=2):
        v = self.table[key]
        assert tuple(v.shape) == tuple(shape) and minval == 0 and maxval == 2
        return A(v)


def reference_ask(pop, lb, ub, pro_c, dis_c, pro_m, dis_m, seed, gen):
    n, d = pop.shape
    dr = philox.variation_draws(n, d, seed, gen)
    rnd = FakeRandom()
    # sbx.py:10  mu_key, beta1_key, beta2_key, beta3_key = random.split(key, 4)
    rnd.table[("x", 0)], rnd.table[("x", 1)] = dr["sbx_mu"], dr["sbx_sign"]
    rnd.table[("x", 2)], rnd.table[("x", 3)] = dr["sbx_keep"], dr["sbx_pair"]
    # pm_mutation.py:9   subkey1, subkey2 = random.split(key)
    rnd.table[("m", 0)], rnd.table[("m", 1)] = dr["pm_site"], dr["pm_mu"]
    ns = {"jnp": jax_shim.jnp, "random": rnd, "jit": jax_shim._jit, "vmap": jax_shim._vmap, "partial": partial,
          "jit_class": lambda c: c}
    jax_shim.lift("operators/crossover/sbx.py", ("simulated_binary",), ns)
    jax_shim.lift("operators/mutation/pm_mutation.py", ("polynomial",), ns)
    crossed = ns["simulated_binary"]("x", A(pop), pro_c, dis_c, 1)
    mutated = ns["polynomial"]("m", crossed, A(np.stack([lb, ub])), pro_m, dis_m)
    return np.clip(np.asarray(mutated), lb, ub).astype(np.float32)        # nsga2.py:85


def main():
    rng = np.random.default_rng(20261108)
    out = {}
    cases = [  # n, d, pro_c, dis_c, pro_m, dis_m, seed, gen, lb, ub
        (64, 7, 1.0, 20.0, 1.0, 20.0, 1, 0, 0.0, 1.0),          # NSGA-II defaults (sbx.py:55, pm_mutation.py:67)
        (1001, 12, 1.0, 20.0, 1.0, 20.0, 7, 3, 0.0, 1.0),       # odd n: the last row passes through
        (400, 5, 0.6, 15.0, 5.0, 10.0, 99, 41, -2.0, 3.0),      # every gene mutates (pro_m = d), some pairs not crossed
        (256, 30, 1.0, 2.0, 3.0, 40.0, 123456, 1000, -1.0, 1.0),
        (2, 1, 1.0, 20.0, 1.0, 20.0, 5, 5, 0.0, 1.0),
    ]
    for i, (n, d, pc, dc, pm, dm, seed, gen, lo, hi) in enumerate(cases):
        lb = np.full(d, lo, np.float32) + (rng.random(d, dtype=np.float32) * 0.1 if i >= 2 else 0)
        ub = np.full(d, hi, np.float32)
        pop = (rng.random((n, d), dtype=np.float32) * (ub - lb) + lb).astype(np.float32)
        if i == 2:
            pop[::7] = ub                                                  # rows on the boundary
            pop[3::11, 0] = lb[0]
        out[f"var{i}_pop"], out[f"var{i}_lb"], out[f"var{i}_ub"] = pop, lb.astype(np.float32), ub
        out[f"var{i}_par"] = np.array([pc, dc, pm, dm], np.float32)
        out[f"var{i}_seed"] = np.array([seed, gen], np.int64)
        out[f"var{i}_out"] = reference_ask(pop, lb.astype(np.float32), ub, pc, dc, pm, dm, seed, gen)
    np.savez_compressed(HERE / "variation_outputs.npz", **out)
    print("wrote", HERE / "variation_outputs.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
