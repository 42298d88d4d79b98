"""Fixtures for SURVEY.md section 8 row f3 (NSGA-III environmental selection), produced by EXECUTING THE
REFERENCE'S OWN SOURCE: the `tell` method of algorithms/mo/nsga3.py:94-210 is lifted out of the class by
ast (read in place from /root/reference) and run on the NumPy-backed jax shim, with the reference's own
non_dominated_sort and cos_dist underneath.  `jax.random.split / randint` are served by the Threefry
restatement in oracle/jax_prng.py (jaxlib itself is absent: that part is restated, not executed).
Run in the build container only:

    python tests/golden/make_golden_nsga3.py
"""
import ast
import sys
import types
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parent.parent))
import jax_shim  # noqa: E402
from oracle import jax_prng  # noqa: E402

ref = jax_shim.install()
A = ref.asarray
jnp, jax = jax_shim.jnp, jax_shim.jax
W = jax_shim._wrapfn

# ---- the few extra jax names nsga3.tell needs ---------------------------------------------------
for name in ("nanmax", "nanargmin", "eye", "isinf", "nansum", "bincount", "isnan"):
    setattr(jnp, name, W(getattr(np, name)))
jnp.nan = np.nan
jnp.bincount = lambda x, length=None: jax_shim._narrow(np.bincount(np.asarray(x).astype(np.int64), minlength=length)[:length])


def _nanargmin(x, axis=None):
    x = np.asarray(x)
    allnan = np.isnan(x).all(axis=axis)
    r = np.nanargmin(np.where(np.expand_dims(allnan, axis) if axis is not None else allnan, 0, x), axis=axis)
    return jax_shim._narrow(np.where(allnan, -1, r).astype(np.int32))      # jnp.nanargmin: -1 for an all-NaN slice


jnp.nanargmin = lambda x, axis=None: _nanargmin(jax_shim._base(x), axis)
jnp.linalg.solve = W(lambda a, b: np.linalg.solve(a, b).astype(np.float32))
jnp.linalg.matrix_rank = W(np.linalg.matrix_rank)


class _Random:
    @staticmethod
    def split(key, num=2):
        return jax_prng.split(key, num)

    @staticmethod
    def randint(key, shape, minval, maxval):
        hi, lo = jax_prng.randint_bits(key)
        return jax_shim._narrow(np.int32(jax_prng.randint_from_bits(hi, lo, int(minval), int(maxval))))


jax.random = _Random


def lifted_tell():
    src = (jax_shim.REF_SRC / "algorithms/mo/nsga3.py").read_text()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "NSGA3")
    fn = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "tell")
    ns = {"jax": jax, "jnp": jnp, "non_dominated_sort": ref.nd.non_dominated_sort, "jit": jax_shim._jit, "vmap": jax_shim._vmap}
    jax_shim.lift("utils/common.py", ("cos_dist",), ns)
    exec(compile(ast.Module([fn], []), "nsga3.py", "exec"), ns)
    return ns["tell"]


class _State:
    def __init__(self, **kw):
        self.__dict__.update(kw)

    def replace(self, **kw):
        d = dict(self.__dict__)
        d.update(kw)
        return _State(**d)


def das_dennis(m, h):
    """Uniform reference directions on the simplex (what sampling.UniformSampling produces)."""
    out = []

    def rec(prefix, left, k):
        if k == m - 1:
            out.append(prefix + [left])
            return
        for i in range(left + 1):
            rec(prefix + [i], left - i, k + 1)
    rec([], h, 0)
    return (np.array(out, np.float32) / np.float32(h)).astype(np.float32)


def main():
    tell = lifted_tell()
    rng = np.random.default_rng(20261109)
    out = {}
    cases = [(30, 3, 6, "random"), (60, 3, 8, "sphere"), (40, 2, 20, "random"), (50, 5, 3, "sphere"), (36, 3, 6, "ties")]
    for i, (pop, m, h, kind) in enumerate(cases):
        refp = das_dennis(m, h)
        n = 2 * pop
        fit = rng.random((n, m), dtype=np.float32) + np.float32(0.05)
        if kind == "sphere":
            fit = (fit / np.linalg.norm(fit, axis=1, keepdims=True) * (1 + 0.3 * rng.random((n, 1), dtype=np.float32))).astype(np.float32)
        elif kind == "ties":
            fit = (rng.integers(1, 6, (n, m)) / 5).astype(np.float32)
        popu = rng.random((n, 4), dtype=np.float32)
        key = jax_prng.prng_key(1000 + i)
        me = types.SimpleNamespace(pop_size=pop, n_objs=m, ref=A(refp))
        st = _State(population=A(popu[:pop]), next_generation=A(popu[pop:]), fitness=A(fit[:pop]), key=key)
        new = tell(me, st, A(fit[pop:]))
        out[f"n3_{i}_pop"], out[f"n3_{i}_fit"], out[f"n3_{i}_ref"] = popu, fit, refp
        out[f"n3_{i}_popsize"] = np.array(pop)
        out[f"n3_{i}_key"] = np.array([int(key[0]), int(key[1])], np.uint32)
        out[f"n3_{i}_opop"] = np.asarray(new.population).astype(np.float32)
        out[f"n3_{i}_ofit"] = np.asarray(new.fitness).astype(np.float32)
    np.savez_compressed(HERE / "nsga3_outputs.npz", **out)
    print("wrote", HERE / "nsga3_outputs.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
