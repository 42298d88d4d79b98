"""NumPy-backed stand-in for the few `jax` names the reference's hot-path files use.

TEST INFRASTRUCTURE ONLY.  JAX is not installed in the build container, so the
reference's own Python sources (read from /root/reference, never copied) cannot
be imported as they are.  This shim provides just enough of `jax`, `jax.numpy`,
`jax.lax` and the `evox` package skeleton for these files to execute under
NumPy with JAX's default dtype rules (x64 disabled: every float result is
float32, every integer result int32):

    src/evox/utils/common.py                     (_dominate, dominate_relation)
    src/evox/operators/selection/non_dominate.py
    src/evox/problems/numerical/{dtlz,lsmop,maf,sphere,griewank,rosenbrock,ackley}.py

`make_golden.py` uses it to produce the fixtures under tests/golden/.  Nothing in
the product path imports this file.
"""
from __future__ import annotations

import ast
import importlib.util
import math
import sys
import types
from pathlib import Path

import numpy as np

REF_SRC = Path("/root/reference/src/evox")


# --------------------------------------------------------------------------- #
# array type with JAX default-dtype behaviour
# --------------------------------------------------------------------------- #
def _narrow(a):
    if isinstance(a, np.ndarray):
        if a.dtype == np.float64:
            a = a.astype(np.float32)
        elif a.dtype == np.int64:
            a = a.astype(np.int32)
        return a.view(JArray)
    if isinstance(a, np.float64):
        return np.float32(a).reshape(()).view(JArray)
    if isinstance(a, np.int64):
        return np.int32(a).reshape(()).view(JArray)
    if isinstance(a, np.generic):
        return np.asarray(a).view(JArray)
    return a


def _base(a):
    return a.view(np.ndarray) if isinstance(a, JArray) else a


class _AtIndexer:
    def __init__(self, arr):
        self.arr = arr

    def __getitem__(self, idx):
        return _AtOp(self.arr, idx)


class _AtOp:
    def __init__(self, arr, idx):
        self.arr, self.idx = arr, idx

    def _idx(self):
        i = self.idx
        if isinstance(i, tuple):
            return tuple(_base(j) for j in i)
        return _base(i)

    def set(self, v):
        out = np.array(_base(self.arr), copy=True)
        out[self._idx()] = np.asarray(_base(v)).astype(out.dtype)
        return _narrow(out)

    def power(self, v):
        out = np.array(_base(self.arr), copy=True)
        cur = out[self._idx()]
        if isinstance(v, int) and out.dtype == np.float32:
            # jnp.power with a Python int exponent lowers to lax.integer_pow:
            # square-and-multiply in float32.
            cur = _integer_pow(cur, v)
        else:
            cur = np.power(cur, v).astype(out.dtype)
        out[self._idx()] = cur
        return _narrow(out)


def _integer_pow(x, y):
    """lax.integer_pow: exponentiation by squaring, float32 roundings kept."""
    assert y >= 0
    acc = None
    base = x.astype(np.float32)
    while y > 0:
        if y & 1:
            acc = base if acc is None else (acc * base).astype(np.float32)
        y >>= 1
        if y > 0:
            base = (base * base).astype(np.float32)
    return np.ones_like(x) if acc is None else acc


class JArray(np.ndarray):
    __array_priority__ = 100

    @property
    def at(self):
        return _AtIndexer(self)

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kw):
        ins = [_base(i) for i in inputs]
        if out is not None:
            kw["out"] = tuple(_base(o) for o in out)
        res = getattr(ufunc, method)(*ins, **kw)
        probe = res[0] if isinstance(res, tuple) else res
        if isinstance(probe, (np.ndarray, np.generic)) and probe.dtype == np.float64:
            # JAX computes in float32: redo with array inputs converted first so
            # there is one rounding, not two.
            ins32 = []
            for i in ins:
                if isinstance(i, (np.ndarray, np.generic)) and i.dtype.kind in "iufb":
                    ins32.append(np.asarray(i).astype(np.float32))
                else:
                    ins32.append(i)
            res = getattr(ufunc, method)(*ins32, **kw)
        if isinstance(res, tuple):
            return tuple(_narrow(r) for r in res)
        return _narrow(res)

    def __pow__(self, y):
        if isinstance(y, int) and self.dtype == np.float32 and y >= 0:
            return _narrow(_integer_pow(_base(self), y))
        return np.power(self, y)

    def __getitem__(self, idx):
        if isinstance(idx, tuple):
            idx = tuple(_base(i) for i in idx)
        else:
            idx = _base(idx)
        r = np.ndarray.__getitem__(_base(self), idx)
        return _narrow(r) if isinstance(r, (np.ndarray, np.generic)) else r

    def __hash__(self):
        return id(self)


def _wrapfn(fn):
    def w(*a, **k):
        a = [_base(x) if isinstance(x, JArray) else
             (type(x)(_base(y) for y in x) if isinstance(x, (list, tuple)) else x) for x in a]
        k = {kk: _base(v) for kk, v in k.items()}
        r = fn(*a, **k)
        if isinstance(r, tuple):
            return tuple(_narrow(x) for x in r)
        if isinstance(r, list):
            return [_narrow(x) for x in r]
        return _narrow(r)
    return w


# --------------------------------------------------------------------------- #
# jax.numpy
# --------------------------------------------------------------------------- #
jnp = types.ModuleType("jax.numpy")
for _n in ("sum", "cos", "sin", "exp", "sqrt", "abs", "max", "min", "maximum", "minimum",
           "tile", "fliplr", "cumprod", "cumsum", "prod", "mean", "floor", "ceil", "stack",
           "hstack", "vstack", "concatenate", "where", "any", "all", "power", "arange",
           "zeros_like", "ones_like", "linspace", "meshgrid", "argmin", "roll", "nanmin",
           "broadcast_to", "matmul", "dot", "copy"):
    setattr(jnp, _n, _wrapfn(getattr(np, _n)))
jnp.pi, jnp.e, jnp.inf, jnp.newaxis = np.pi, np.e, np.inf, np.newaxis
jnp.bool_, jnp.int32, jnp.float32 = np.bool_, np.int32, np.float32
jnp.ndarray = np.ndarray


def _shape(x):
    return np.shape(x)


def _asarray(x, dtype=None):
    return _narrow(np.asarray(_base(x), dtype=dtype))


def _zeros(shape, dtype=np.float32):
    return _narrow(np.zeros(shape, dtype=dtype))


def _ones(shape, dtype=np.float32):
    return _narrow(np.ones(shape, dtype=dtype))


def _lexsort(keys):
    # stable; NaN last; -0 == +0 (matches jnp.lexsort's float canonicalisation)
    return _narrow(np.lexsort(tuple(_base(k) for k in keys)).astype(np.int32))


def _argsort(x, axis=-1):
    return _narrow(np.argsort(_base(x), axis=axis, kind="stable").astype(np.int32))


def _unique(ar, axis=None, size=None, return_index=False, return_counts=False):
    """jnp.unique(..., size=) semantics of jax 0.4.x: the value pad is the
    minimum row, the index pad is the index of the first sorted row, count pad 0."""
    a = _base(ar)
    vals, idx, cnt = np.unique(a, axis=axis, return_index=True, return_counts=True)
    if size is not None and vals.shape[0] < size:
        pad = size - vals.shape[0]
        vals = np.concatenate([vals, np.repeat(vals[:1], pad, axis=0)], axis=0)
        idx = np.concatenate([idx, np.repeat(idx[:1], pad)])
        cnt = np.concatenate([cnt, np.zeros(pad, cnt.dtype)])
    out = [_narrow(vals)]
    if return_index:
        out.append(_narrow(idx))
    if return_counts:
        out.append(_narrow(cnt))
    return tuple(out) if len(out) > 1 else out[0]


class _CClass:
    def __getitem__(self, key):
        key = tuple(_base(k) for k in key) if isinstance(key, tuple) else _base(key)
        return _narrow(np.c_[key])


class _RClass:
    def __getitem__(self, key):
        key = tuple(_base(k) for k in key) if isinstance(key, tuple) else _base(key)
        return _narrow(np.r_[key])


jnp.shape = _shape
jnp.asarray = _asarray
jnp.array = _asarray
jnp.zeros = _zeros
jnp.ones = _ones
jnp.empty = _zeros           # XLA materialises jnp.empty as zeros
jnp.lexsort = _lexsort
jnp.argsort = _argsort
jnp.sort = _wrapfn(lambda x, axis=-1: np.sort(x, axis=axis, kind="stable"))
jnp.unique = _unique
jnp.c_ = _CClass()
jnp.r_ = _RClass()
jnp.linalg = types.SimpleNamespace(norm=_wrapfn(np.linalg.norm))


# --------------------------------------------------------------------------- #
# jax, jax.lax
# --------------------------------------------------------------------------- #
def _jit(fun=None, **_kw):
    if fun is None:
        return lambda f: f
    return fun


def _vmap(fun, in_axes=0, out_axes=0):
    def mapped(*args):
        axes = in_axes if isinstance(in_axes, (tuple, list)) else (in_axes,) * len(args)
        n = None
        for a, ax in zip(args, axes):
            if ax is not None:
                n = np.shape(a)[ax]
                break
        outs = []
        for i in range(n):
            call = [a if ax is None else np.take(_base(a), i, axis=ax).view(JArray)
                    for a, ax in zip(args, axes)]
            outs.append(fun(*call))
        if n == 0:
            probe = fun(*[a if ax is None else np.zeros(np.delete(np.shape(a), ax), np.asarray(a).dtype).view(JArray)
                          for a, ax in zip(args, axes)])
            shp = list(np.shape(probe))
            shp.insert(out_axes, 0)
            return _narrow(np.zeros(shp, np.asarray(probe).dtype))
        return _narrow(np.stack([np.asarray(_base(o)) for o in outs], axis=out_axes))
    return mapped


def _while_loop(cond, body, init):
    st = init
    while bool(cond(st)):
        st = body(st)
    return st


def _scan(f, init, xs):
    carry = init
    n = len(xs[0]) if isinstance(xs, tuple) else len(xs)
    ys = []
    for i in range(n):
        x = tuple(a[i] for a in xs) if isinstance(xs, tuple) else xs[i]
        carry, y = f(carry, x)
        ys.append(y)
    return carry, None if all(y is None for y in ys) else ys


def _cond(pred, tfn, ffn, *ops):
    return tfn(*ops) if bool(pred) else ffn(*ops)


def _pure_callback(fn, _shape, *args):
    return _narrow(fn(*[np.asarray(_base(a)) for a in args]))


jax = types.ModuleType("jax")
lax = types.ModuleType("jax.lax")
lax.while_loop, lax.scan, lax.cond = _while_loop, _scan, _cond
jax.numpy, jax.lax = jnp, lax
jax.jit, jax.vmap, jax.pure_callback = _jit, _vmap, _pure_callback
jax.Array = np.ndarray
jax.default_backend = lambda: "cpu"
jax.ShapeDtypeStruct = lambda shape, dtype=None: (shape, dtype)


def install():
    """Put the shim modules and a skeletal `evox` package into sys.modules and
    load the reference hot-path files from /root/reference on top of them."""
    sys.modules["jax"] = jax
    sys.modules["jax.numpy"] = jnp
    sys.modules["jax.lax"] = lax

    class Problem:
        def __init__(self, *a, **k):
            pass

        def init(self, key=None):
            return self.setup(key)

        def setup(self, key):
            return State(key=key)

    class State(dict):
        def __init__(self, **kw):
            super().__init__(**kw)

    class _Sampling:
        def __init__(self, *a, **k):
            pass

    evox = types.ModuleType("evox")
    evox.__path__ = []
    evox.Problem, evox.State = Problem, State
    evox.jit_class = lambda cls: cls
    sys.modules["evox"] = evox

    def pkg(name):
        m = types.ModuleType(name)
        m.__path__ = []
        sys.modules[name] = m
        return m

    utils = pkg("evox.utils")
    ops = pkg("evox.operators")
    samp = pkg("evox.operators.sampling")
    sel = pkg("evox.operators.selection")
    probs = pkg("evox.problems")
    num = pkg("evox.problems.numerical")
    samp.UniformSampling = samp.GridSampling = _Sampling

    # evox.utils.dominate_relation: common.py imports optax etc. at module top,
    # so lift only the two functions out by AST and run them on the shim.
    src = (REF_SRC / "utils/common.py").read_text()
    tree = ast.parse(src)
    ns = {"jax": jax, "jnp": jnp, "jit": _jit, "vmap": _vmap}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in ("_dominate", "dominate_relation"):
            exec(compile(ast.Module([node], []), "common.py", "exec"), ns)
    utils.dominate_relation = ns["dominate_relation"]
    evox.utils = utils

    def load(modname, rel):
        spec = importlib.util.spec_from_file_location(modname, REF_SRC / rel)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
        return mod

    for f in ("sphere", "griewank", "rosenbrock", "ackley"):
        m = load(f"evox.problems.numerical.{f}", f"problems/numerical/{f}.py")
        setattr(num, f"{f}_func", getattr(m, f"{f}_func"))
        setattr(num, f.capitalize(), getattr(m, f.capitalize()))
    nd = load("evox.operators.selection.non_dominate", "operators/selection/non_dominate.py")
    for n in ("non_dominated_sort", "crowding_distance", "crowding_distance_sort",
              "non_dominate", "NonDominate", "host_rank_from_domination_matrix"):
        setattr(ops, n, getattr(nd, n))
    dtlz = load("evox.problems.numerical.dtlz", "problems/numerical/dtlz.py")
    lsmop = load("evox.problems.numerical.lsmop", "problems/numerical/lsmop.py")

    # maf.py: MaF2..15 pull in more of jax than the shim has; lift MaF + MaF1.
    tree = ast.parse((REF_SRC / "problems/numerical/maf.py").read_text())
    ns = {"jax": jax, "jnp": jnp, "evox": evox, "Problem": Problem, "State": State,
          "UniformSampling": _Sampling, "math": math}
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name in ("MaF", "MaF1"):
            exec(compile(ast.Module([node], []), "maf.py", "exec"), ns)
    maf = types.SimpleNamespace(MaF1=ns["MaF1"])
    return types.SimpleNamespace(utils=utils, nd=nd, dtlz=dtlz, lsmop=lsmop, maf=maf,
                                 asarray=_asarray)


def lift(rel, names, ns):
    """Execute only the named top-level functions / classes of a reference source file in `ns`
    (for files whose module top imports more of jax / evox than the shim provides)."""
    tree = ast.parse((REF_SRC / rel).read_text())
    for node in tree.body:
        if isinstance(node, (ast.FunctionDef, ast.ClassDef)) and node.name in names:
            exec(compile(ast.Module([node], []), rel, "exec"), ns)
    return ns


def lift_spea2():
    """cal_fitness and _truncation of algorithms/mo/spea2.py:21-67 with the utils they call."""
    from functools import partial
    ns = {"jax": jax, "jnp": jnp, "jit": _jit, "vmap": _vmap, "partial": partial}
    lift("utils/common.py", ("pairwise_func", "euclidean_dist", "pairwise_euclidean_dist", "_dominate",
                             "dominate_relation"), ns)
    lift("algorithms/mo/spea2.py", ("cal_fitness", "_truncation"), ns)
    return types.SimpleNamespace(cal_fitness=ns["cal_fitness"], truncation=ns["_truncation"])

