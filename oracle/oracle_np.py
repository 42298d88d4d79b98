"""CPU restatement (NumPy, float32) of EvoX's multi-objective environmental-selection path.

TEST INFRASTRUCTURE, NOT PRODUCT.  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this package; the
product path (evox_b200/) never does and fails loudly without its CUDA library.

Every function cites the reference lines it follows (paths relative to
/root/reference/).  The reference executes these through jaxlib/XLA, a third-party
dependency that is absent here and unpinned (pyproject.toml:33-34, `jax>=0.4.16`,
dev shell ~0.4.20), so this is a restatement of the reference's *source text*.

Pinning (tests/test_oracle_*.py):
  * every explicit-input known-answer test of the reference for this path
    (tests/test_non_dominated_sort.py:8-28, tests/test_crowding_distance.py:7-21 and
    :127-146, tests/test_monitors.py:30-48, tests/test_classic_problems.py:46-54);
  * the JAX-PRNG-input goldens (tests/test_crowding_distance.py:24-124,
    tests/test_lsmop.py:19-124, tests/test_maf.py:49-57) with the inputs regenerated
    by oracle/jax_prng.py;
  * fixtures under tests/golden/ produced by executing the reference's own .py
    files on a NumPy-backed jax shim (tests/golden/make_golden.py).
"""
from itertools import cycle

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------- #
# dominance                                                 utils/common.py:88-97
# --------------------------------------------------------------------------- #
def dominate_relation(x, y):
    """A[i, j] = all_k(x[i,k] <= y[j,k]) and any_k(x[i,k] < y[j,k]).

    src/evox/utils/common.py:88-91 (`_dominate`), :94-97 (nested vmap).  IEEE
    compares: a NaN anywhere in a row makes it neither dominate nor be dominated.
    """
    x = np.asarray(x)
    y = np.asarray(y)
    nx, ny = x.shape[0], y.shape[0]
    m = x.shape[1] if x.ndim == 2 else 0
    out = np.empty((nx, ny), dtype=bool)
    step = max(1, (1 << 22) // max(1, ny * max(1, m)))
    with np.errstate(invalid="ignore"):
        for s in range(0, nx, step):
            xs = x[s:s + step, None, :]
            out[s:s + step] = (xs <= y[None, :, :]).all(-1) & (xs < y[None, :, :]).any(-1)
    return out


def rank_from_domination_matrix(dom, count):
    """Front peeling on a dense matrix.

    src/evox/operators/selection/non_dominate.py:13-29 (the reference's NumPy
    `host` variant); the jitted variants :74-113 compute the same thing: a row
    enters front k when its dominator count reaches zero, then every row it
    dominates loses one, and the row itself is pushed to -1 so it never returns.
    """
    count = np.array(count, dtype=np.int64, copy=True)
    n = dom.shape[0]
    rank = np.zeros(n, dtype=np.int32)          # jitted variants start from zeros (:74)
    k = 0
    front = count == 0
    while front.any():
        rank[front] = k
        count -= dom[front, :].sum(axis=0)
        count -= front
        k += 1
        front = count == 0
    return rank


def non_dominated_sort(x):
    """src/evox/operators/selection/non_dominate.py:32-115: matrix (:62), column
    sums (:63), peel (:74-113).  Returns int32 ranks starting at 0."""
    x = np.asarray(x)
    dom = dominate_relation(x, x)
    return rank_from_domination_matrix(dom, dom.sum(axis=0))


# --------------------------------------------------------------------------- #
# crowding distance                                      non_dominate.py:118-180
# --------------------------------------------------------------------------- #
def crowding_distance(costs, mask=None):
    """src/evox/operators/selection/non_dominate.py:118-158.

    Per objective: stable lexsort by (valid first, then cost) (:147); sorted
    costs; range = c[nv-1]-c[0] (:149); every interior *position* of the full-length
    order gets (c[i+1]-c[i-1])/range (:151); positions 0 and nv-1 get +inf
    (:152-153); masked rows become -inf (:155); objectives are summed (:158).
    nv-1 == -1 wraps to the last position exactly as JAX indexing does.
    """
    costs = np.asarray(costs, dtype=F32)
    n = costs.shape[0]
    if mask is None:
        nv = n
        mask = np.ones(n, dtype=bool)
    else:
        mask = np.asarray(mask, dtype=bool)
        nv = int(mask.sum())
    total = np.zeros(n, dtype=F32)
    if n == 0:
        return total
    with np.errstate(invalid="ignore", divide="ignore"):
        for k in range(costs.shape[1]):
            c = costs[:, k]
            order = np.lexsort((c, ~mask))
            cs = c[order]
            rng = F32(cs[nv - 1] - cs[0])
            d = np.zeros(n, dtype=F32)                      # jnp.empty -> zeros
            d[order[1:-1]] = (cs[2:] - cs[:-2]) / rng
            d[order[0]] = np.inf
            d[order[nv - 1]] = np.inf
            d = np.where(mask, d, F32(-np.inf)).astype(F32)
            total = (total + d).astype(F32) if k else d
    return total


def crowding_distance_sort(x, mask=None):
    """src/evox/operators/selection/non_dominate.py:161-180: stable ascending argsort."""
    return np.argsort(crowding_distance(x, mask), kind="stable").astype(np.int32)


# --------------------------------------------------------------------------- #
# survivor selection                                     non_dominate.py:183-232
# --------------------------------------------------------------------------- #
def non_dominate(population, fitness, topk, deduplicate=False):
    """src/evox/operators/selection/non_dominate.py:183-222.

    Keeps the reference's quirk: :216 stores the crowding *argsort permutation* and
    :218 uses its negation as the secondary sort key.  The dedupe branch (:188-208)
    follows jnp.unique(size=N) of jax 0.4.x (index pad = first sorted row); that
    padding rule belongs to jaxlib and is not pinned by any reference test.
    """
    population = np.asarray(population)
    fitness = np.asarray(fitness, dtype=F32)
    fitness_bak = fitness
    if deduplicate:
        n = population.shape[0]
        _, uidx, ucnt = np.unique(population, axis=0, return_index=True, return_counts=True)
        count = uidx.shape[0]
        pad = n - count
        uidx = np.concatenate([uidx, np.repeat(uidx[:1], pad)])
        population = population[uidx]
        fitness = fitness[uidx]
        fitness_bak = fitness
        fitness = np.where((np.arange(n) < count)[:, None], fitness, F32(np.inf)).astype(F32)
    rank = non_dominated_sort(fitness)
    order = np.argsort(rank, kind="stable")
    worst = rank[order[topk - 1]]
    mask = rank == worst
    cd_perm = crowding_distance_sort(fitness, mask)
    combined = np.lexsort((-cd_perm, rank))[:topk]
    return population[combined], fitness_bak[combined]


# --------------------------------------------------------------------------- #
# DTLZ                                        problems/numerical/dtlz.py:36-322
# --------------------------------------------------------------------------- #
def _shape_linear(scale, X, m):
    """(scale) * fliplr(cumprod([1, x_0..x_{m-2}])) * [1, 1-x_{m-2}, ..., 1-x_0]   dtlz.py:62-67"""
    n = X.shape[0]
    cp = np.cumprod(np.c_[np.ones((n, 1), F32), X[:, :m - 1]], axis=1, dtype=F32)
    return (scale * np.fliplr(cp) * np.c_[np.ones((n, 1), F32), 1 - X[:, m - 2::-1]]).astype(F32)


def _shape_sphere(scale, X, m, clamp=True):
    """dtlz.py:87-101: cos/sin(x*pi/2) staircase, cos clamped at 0 in DTLZ2-6 (not in LSMOP5-8)."""
    n = X.shape[0]
    cs = np.cos(X[:, :m - 1] * np.pi / 2).astype(F32)
    if clamp:
        cs = np.maximum(cs, 0)
    cp = np.cumprod(np.c_[np.ones((n, 1), F32), cs], axis=1, dtype=F32)
    sn = np.sin(X[:, m - 2::-1] * np.pi / 2).astype(F32)
    return (scale * np.fliplr(cp) * np.c_[np.ones((n, 1), F32), sn]).astype(F32)


def _g_multimodal(X, m):
    """100*(k + sum((x-.5)^2 - cos(20*pi*(x-.5))))   dtlz.py:52-61, :118-131"""
    d = X.shape[1]
    t = X[:, m - 1:] - F32(0.5)
    s = np.sum(t * t - np.cos(20 * np.pi * t), axis=1, keepdims=True, dtype=F32)
    return (100 * (d - m + 1 + s)).astype(F32)


def _integer_pow_f32(x, y):
    """lax.integer_pow (what `X.at[...].power(100)` lowers to): square-and-multiply in f32."""
    acc, base = None, x.astype(F32)
    while y > 0:
        if y & 1:
            acc = base if acc is None else (acc * base).astype(F32)
        y >>= 1
        if y:
            base = (base * base).astype(F32)
    return acc


def dtlz(which, X, m):
    """DTLZ1..7 evaluate.  dtlz.py:48-69, 84-103, 115-145, 152-172, 188-208, 244-265, 303-322."""
    X = np.array(X, dtype=F32, copy=True)
    n, d = X.shape
    with np.errstate(all="ignore"):
        if which == 1:
            g = _g_multimodal(X, m)
            return _shape_linear(F32(0.5) * (1 + g), X, m)
        if which == 2:
            g = np.sum((X[:, m - 1:] - F32(0.5)) ** 2, axis=1, keepdims=True, dtype=F32)
            return _shape_sphere(1 + g, X, m)
        if which == 3:
            g = _g_multimodal(X, m)
            return _shape_sphere(1 + g, X, m)
        if which == 4:
            X[:, :m - 1] = _integer_pow_f32(X[:, :m - 1], 100)
            g = np.sum((X[:, m - 1:] - F32(0.5)) ** 2, axis=1, keepdims=True, dtype=F32)
            return _shape_sphere(1 + g, X, m)
        if which in (5, 6):
            if which == 5:
                g = np.sum((X[:, m - 1:] - F32(0.5)) ** 2, axis=1, keepdims=True, dtype=F32)
            else:
                g = np.sum(np.power(X[:, m - 1:], F32(0.1)), axis=1, keepdims=True, dtype=F32)
            X[:, 1:m - 1] = (1 + 2 * g * X[:, 1:m - 1]) / (2 + 2 * g)
            return _shape_sphere(1 + g, X, m)
        if which == 7:
            g = 1 + 9 * np.mean(X[:, m - 1:], axis=1, keepdims=True, dtype=F32)
            f = np.zeros((n, m), F32)
            f[:, :m - 1] = X[:, :m - 1]
            h = m - np.sum(f[:, :m - 1] / (1 + g) * (1 + np.sin(3 * np.pi * f[:, :m - 1])),
                           axis=1, keepdims=True, dtype=F32)
            f[:, m - 1:] = (1 + g) * h
            return f
    raise ValueError(which)


# --------------------------------------------------------------------------- #
# MaF1                                         problems/numerical/maf.py:95-108
# --------------------------------------------------------------------------- #
def maf1(X, m):
    X = np.asarray(X, dtype=F32)
    g = np.sum((X[:, m - 1:] - F32(0.5)) ** 2, axis=1, dtype=F32).reshape(-1, 1)
    rg = 1 + g
    return (rg - _shape_linear(rg, X, m)).astype(F32)


# --------------------------------------------------------------------------- #
# LSMOP                                       problems/numerical/lsmop.py:17-454
# --------------------------------------------------------------------------- #
def lsmop_groups(d, m, nk=5):
    """Sub-group lengths and offsets.  lsmop.py:46-54: logistic map c_i, then
    floor(c/sum(c)*d/nk) evaluated in float32, cumulative offsets of sublen*nk."""
    c = [3.8 * 0.1 * (1 - 0.1)]
    for _ in range(1, m):
        c.append(3.8 * c[-1] * (1 - c[-1]))
    c = np.asarray(c, dtype=F32)
    sub = np.floor(c / np.sum(c, dtype=F32) * d / nk)
    sublen = tuple(int(v) for v in sub)
    off = tuple(int(v) for v in np.r_[0, np.cumsum(sub * nk)])
    return sublen, off


def _sphere(x):                      # sphere.py:5-6
    return np.sum(x * x, axis=-1, dtype=F32)


def _griewank(x):                    # griewank.py:6-12
    k = np.sqrt(np.arange(1, x.shape[1] + 1, dtype=F32))
    return (F32(1 / 4000) * np.sum(x * x, axis=1, dtype=F32)
            - np.prod(np.cos(x / k), axis=1, dtype=F32) + 1).astype(F32)


def _rosenbrock(x):                  # rosenbrock.py:6-12
    a, b = x[:, :-1], x[:, 1:]
    return np.sum(100 * (b - a * a) ** 2 + (a - 1) ** 2, axis=1, dtype=F32)


def _ackley(x, a=20, b=0.2, c=2 * np.pi):   # ackley.py:7-13, defaults lsmop.py:59
    return (-a * np.exp(-b * np.sqrt(np.mean(x * x, axis=1, dtype=F32)))
            - np.exp(np.mean(np.cos(c * x), axis=1, dtype=F32)) + a + np.e).astype(F32)


def _schwefel(x):                    # lsmop.py:76-77
    return np.max(np.abs(x), axis=-1)


def _rastrigin(x):                   # lsmop.py:84-86
    return np.sum(x * x - 10 * np.cos(2 * np.pi * x) + 10, axis=1, dtype=F32)


_LSMOP_INNER = {          # lsmop.py:136, 161, 188, 215, 240, 277, 309, 348, 386
    1: (_sphere,), 2: (_griewank, _schwefel), 3: (_rastrigin, _rosenbrock),
    4: (_ackley, _griewank), 5: (_sphere,), 6: (_rosenbrock, _schwefel),
    7: (_ackley, _rosenbrock), 8: (_griewank, _sphere), 9: (_sphere, _ackley),
}


def lsmop(which, X, m, sublen=None, off=None, nk=5):
    """LSMOP1..9 evaluate (lsmop.py:127-141 ... 400-426) with `_calc_g` (:88-111).

    `sublen`/`off` default to the groups of d = X.shape[1]; pass the constructor's
    to reproduce the reference's d=None quirk (groups sized for d=100*m)."""
    X = np.array(X, dtype=F32, copy=True)
    n, d = X.shape
    if sublen is None:
        sublen, off = lsmop_groups(d, m, nk)
    with np.errstate(all="ignore"):
        j = np.arange(m, d + 1, dtype=F32) / F32(d)
        w = 1 + j if which <= 4 else 1 + np.cos(j * np.pi / 2).astype(F32)
        X[:, m - 1:d] = w * X[:, m - 1:d] - X[:, :1] * 10
        g = []
        for o, sl, fn in zip(off, sublen, cycle(_LSMOP_INNER[which])):
            acc = None
            for b in range(nk):
                s = o + m - 1 + b * sl
                r = fn(X[:, s:s + sl]).astype(F32)
                acc = r if acc is None else (acc + r).astype(F32)
            g.append(acc)
        g = np.stack(g, axis=1)
        g = (g / np.asarray(sublen, dtype=F32)[None, :] / nk).astype(F32)
        if which <= 4:
            return _shape_linear(1 + g, X, m)
        if which <= 8:
            return _shape_sphere(1 + g + np.c_[g[:, 1:], np.zeros((n, 1), F32)], X, m, clamp=False)
        gs = 1 + np.sum(g, axis=1, keepdims=True, dtype=F32)
        f = np.zeros((n, m), F32)
        f[:, :m - 1] = X[:, :m - 1]
        h = m - np.sum(f[:, :m - 1] / (1 + gs) * (1 + np.sin(3 * np.pi * f[:, :m - 1])),
                       axis=1, keepdims=True, dtype=F32)
        f[:, m - 1:m] = (1 + gs) * h
        return f


# --------------------------------------------------------------------------- #
# SPEA2 fitness and truncation   algorithms/mo/spea2.py
# --------------------------------------------------------------------------- #
def pairwise_euclidean_dist(x, y):
    """d[i, j] = || x[i] - y[j] ||_2 in float32.  src/evox/utils/common.py:33-56
    (`pairwise_func` over `euclidean_dist` = `jnp.linalg.norm(x - y)`)."""
    x = np.asarray(x, F32)
    y = np.asarray(y, F32)
    diff = (x[:, None, :] - y[None, :, :]).astype(F32)
    return np.sqrt((diff * diff).astype(F32).sum(-1, dtype=F32)).astype(F32)


def spea2_cal_fitness(obj):
    """SPEA2 fitness = density + raw fitness.  src/evox/algorithms/mo/spea2.py:21-37.

    strength s[i] = rows i dominates (:25); raw r[j] = sum of s over j's dominators (:27-29);
    density d = 1 / (distance to the k-th nearest other row + 2) with k = floor(sqrt(6)) = 2, i.e.
    sorted-distance index 1 once the diagonal is +inf (:31-36)."""
    obj = np.asarray(obj, F32)
    n = obj.shape[0]
    dom = dominate_relation(obj, obj)
    s = dom.sum(1).astype(np.int32)
    r = (dom * s[:, None]).sum(0).astype(np.int32)
    dis = pairwise_euclidean_dist(obj, obj)
    dis[np.arange(n), np.arange(n)] = np.inf
    dis = np.sort(dis, axis=1)
    kth = min(int(np.floor(np.sqrt(6))) - 1, n - 1)       # n = 1: XLA clamps the out-of-range column index
    d = (F32(1) / (dis[:, kth] + F32(2))).astype(F32)
    return (d + r.astype(F32)).astype(F32)


def spea2_truncation(obj, k, mask):
    """Archive truncation: k times remove the valid row whose nearest valid neighbour is closest
    (ties: lowest index, `jnp.argsort` is stable); returns the rows kept.
    src/evox/algorithms/mo/spea2.py:40-67.  As in the reference, k must not exceed the number of valid
    rows minus one (its while_loop would not terminate otherwise)."""
    obj = np.asarray(obj, F32)
    mask = np.asarray(mask, bool)
    n = obj.shape[0]
    dis = pairwise_euclidean_dist(obj, obj)
    dis[np.arange(n), np.arange(n)] = np.inf
    dis = np.where(mask[:, None] & mask[None, :], dis, F32(np.inf)).astype(F32)
    keep = np.ones(n, bool)
    assert k <= max(0, int(mask.sum()) - 1), "spea2_truncation: k exceeds what the reference's loop can remove"
    for _ in range(int(k)):
        idx = int(np.argsort(dis.min(axis=1), kind="stable")[0])
        keep[idx] = False
        dis[idx, :] = np.inf
        dis[:, idx] = np.inf
    return keep & mask


# --------------------------------------------------------------------------- #
# NSGA-II variation      operators/crossover/sbx.py:8-40,
#                                      operators/mutation/pm_mutation.py:7-54, algorithms/mo/nsga2.py:85
# --------------------------------------------------------------------------- #
def simulated_binary(x, mu, sign, keep, pair, pro_c, dis_c):
    """sbx.py:8-40, type 1, with the four random arrays passed in instead of drawn:
    mu = uniform(mu_key) [pairs, d]; sign = randint(beta1_key, 0, 2); keep = uniform(beta2_key);
    pair = uniform(beta3_key) [pairs, 1]."""
    x = np.asarray(x, F32)
    n = x.shape[0]
    p1, p2 = x[: n // 2], x[n // 2: n // 2 * 2]
    with np.errstate(invalid="ignore", divide="ignore"):
        beta = np.where(mu <= 0.5, np.power(F32(2) * mu, F32(1 / (dis_c + 1))), F32(0)).astype(F32)      # :16
        beta = np.where(mu > 0.5, np.power(F32(2) - F32(2) * mu, F32(-1 / (dis_c + 1))), beta).astype(F32)   # :17
    beta = (beta * np.where(sign & 1, F32(-1), F32(1))).astype(F32)                                      # :18-20
    beta = np.where(keep < 0.5, F32(1), beta)                                                            # :21
    beta = np.where(np.tile(pair > F32(pro_c), (1, x.shape[1])), F32(1), beta).astype(F32)              # :22-24
    mid, dev = (p1 + p2) / F32(2), beta * (p1 - p2) / F32(2)
    out = np.vstack((mid + dev, mid - dev)).astype(F32)                                                  # :26-34
    if n % 2:
        out = np.r_[out, x[-1:]]                                                                         # :35-36
    return out


def polynomial(x, lb, ub, site_u, mu, pro_m, dis_m):
    """pm_mutation.py:7-54 with site_u = uniform(subkey1), mu = uniform(subkey2), both [2*(n//2), d]."""
    x = np.asarray(x, F32)
    pop = x if x.shape[0] == 1 else x[: (x.shape[0] // 2) * 2]
    n, d = pop.shape
    site = site_u < F32(pro_m / d)                                                                       # :16
    lower, upper = np.tile(np.asarray(lb, F32), (n, 1)), np.tile(np.asarray(ub, F32), (n, 1))
    pop = np.maximum(np.minimum(pop, upper), lower)                                                      # :21
    e1, e = F32(dis_m + 1.0), F32(1.0 / (dis_m + 1))
    one, two, half = F32(1), F32(2), F32(0.5)
    with np.errstate(invalid="ignore", divide="ignore"):
        temp = site & (mu <= 0.5)                                                                        # :19
        norm = np.where(temp, (pop - lower) / (upper - lower), F32(0)).astype(F32)
        pop = np.where(temp, pop + (upper - lower) * (np.power(two * mu + (one - two * mu) * np.power(one - norm, e1), e) - one),
                       pop).astype(F32)                                                                  # :23-35
        temp = site & (mu > 0.5)                                                                         # :37
        norm = np.where(temp, (upper - pop) / (upper - lower), F32(0)).astype(F32)
        pop = np.where(temp, pop + (upper - lower) * (one - np.power(two * (one - mu) + two * (mu - half) * np.power(one - norm, e1), e)),
                       pop).astype(F32)                                                                  # :38-51
    if x.shape[0] % 2:
        pop = np.r_[pop, x[-1:]]                                                                         # :52-53
    return pop


def nsga2_variation(pop, lb, ub, pro_c, dis_c, pro_m, dis_m, seed, generation):
    """nsga2.py:78-85 (`ask`: crossover -> mutation -> clip) on the uniforms csrc/variation.cu draws
    (oracle/philox.py) instead of jax.random's."""
    from . import philox
    pop = np.asarray(pop, F32)
    n, d = pop.shape
    if n < 2:
        return np.clip(pop, lb, ub).astype(F32)
    dr = philox.variation_draws(n, d, seed, generation)
    c = simulated_binary(pop, dr["sbx_mu"], dr["sbx_sign"], dr["sbx_keep"], dr["sbx_pair"], pro_c, dis_c)
    c = polynomial(c, lb, ub, dr["pm_site"], dr["pm_mu"], pro_m, dis_m)
    return np.clip(c, np.asarray(lb, F32), np.asarray(ub, F32)).astype(F32)


# --------------------------------------------------------------------------- #
# NSGA-III environmental selection         algorithms/mo/nsga3.py:97-205
# --------------------------------------------------------------------------- #
def cos_dist(x, y):
    """utils/common.py:75-80."""
    x = np.asarray(x, F32)
    y = np.asarray(y, F32)
    with np.errstate(invalid="ignore", divide="ignore"):
        xn = (x / np.linalg.norm(x, axis=-1, keepdims=True)).astype(F32)
        yn = (y / np.linalg.norm(y, axis=-1, keepdims=True)).astype(F32)
    return np.dot(xn, yn.T).astype(F32)


def nsga3_normalize(merged_fitness, rank, last_rank):
    """nsga3.py:103-139: NaN out the rows past the last front, shift by the ideal point, find the
    extreme points, intercepts of their hyperplane (or the worst point when they are degenerate)."""
    import warnings
    fit = np.asarray(merged_fitness, F32)
    m = fit.shape[1]
    with np.errstate(invalid="ignore", divide="ignore"), warnings.catch_warnings():
        warnings.simplefilter("ignore")                       # all-NaN rows are part of the algorithm
        ranked = np.where((rank <= last_rank)[:, None], fit, F32(np.nan)).astype(F32)          # :103-107
        ideal = np.nanmin(ranked, axis=0)                                                     # :110
        ranked = (ranked - ideal).astype(F32)
        weight = (np.eye(m, dtype=F32) + F32(1e-6)).astype(F32)                               # :112
        extreme_ind = np.array([np.nanargmin(np.nanmax((ranked / weight[i]).astype(F32), axis=1)) for i in range(m)])   # :114-117
        extreme = ranked[extreme_ind]
        if np.linalg.matrix_rank(extreme) == m:                                               # :133
            plane = np.linalg.solve(extreme, np.ones(m, F32)).astype(F32)                     # :123
            nadir = (F32(1) / plane).astype(F32)
        else:
            nadir = np.nanmax(ranked, axis=0)                                                 # :130
        return (ranked / nadir).astype(F32)                                                   # :139


def nsga3_associate(normalized, ref):
    """nsga3.py:140-147: perpendicular distance to every reference direction, nearest one per row.
    Rows of NaN (past the last front) get group len(ref) and distance NaN."""
    ref = np.asarray(ref, F32)
    with np.errstate(invalid="ignore", divide="ignore"):
        c = cos_dist(normalized, ref)
        dist = (np.linalg.norm(normalized, axis=-1, keepdims=True).astype(F32) * np.sqrt(F32(1) - (c * c).astype(F32))).astype(F32)
    allnan = np.isnan(dist).all(axis=1)
    gid = np.where(allnan, len(ref), np.nanargmin(np.where(allnan[:, None], F32(0), dist), axis=1)).astype(np.int32)
    with np.errstate(invalid="ignore"), __import__("warnings").catch_warnings():
        __import__("warnings").simplefilter("ignore")
        gdist = np.nanmin(dist, axis=1).astype(F32)
    return gid, gdist


def nsga3_niche(rank, last_rank, group_id, group_dist, nref, pop_size, rbits):
    """nsga3.py:148-205: niche-preserving fill of the last front.  rbits[i] = (higher, lower) words that
    `jax.random.randint(keys[i], ...)` would consume (oracle/jax_prng.randint_bits).  Returns the indices
    of the survivors, ascending (:202-204)."""
    from . import jax_prng
    rank = np.array(rank, np.int64, copy=True)
    n = rank.shape[0]
    last = rank == last_rank
    rho = np.bincount(np.where(rank < last_rank, group_id, nref), minlength=nref + 1)[:nref].astype(np.float64)
    rho_last = np.bincount(np.where(last, group_id, nref), minlength=nref + 1)[:nref].astype(np.int64)
    gid = np.where(last, group_id.astype(np.float64), np.inf)
    gdist = np.where(last, group_dist.astype(np.float64), np.inf)
    selected = int(rho.sum())
    rho = np.where(rho_last == 0, np.inf, rho)
    while selected < pop_size:
        group = int(np.argmin(rho))
        cand = np.where(gid == group, gdist, np.inf)
        if rho[group] == 0 or rho_last[group] == 1:
            c = int(np.argmin(cand))
        else:
            order = np.sort(np.where(np.isinf(cand), np.inf, np.arange(n)))
            hi, lo = rbits[selected]
            c = int(order[jax_prng.randint_from_bits(hi, lo, 0, int(rho_last[group]))])
        rank[c] = last_rank - 1
        gid[c] = np.nan
        rho_last[group] -= 1
        if rho_last[group] == 0:
            rho[group] = np.inf
        else:
            rho[group] += 1
        selected += 1
    return np.nonzero(rank < last_rank)[0][:pop_size].astype(np.int32)


def nsga3_tell(merged_pop, merged_fitness, ref, pop_size, key):
    """NSGA3.tell, nsga3.py:97-210, on an already merged population.  `key` = (hi, lo) words of the
    state's PRNG key.  Returns (population, fitness, intermediates)."""
    from . import jax_prng
    fit = np.asarray(merged_fitness, F32)
    rank = non_dominated_sort(fit)
    order = np.argsort(rank, kind="stable")
    last_rank = int(rank[order[pop_size]])                                                    # :101 (index pop_size)
    normalized = nsga3_normalize(fit, rank, last_rank)
    gid, gdist = nsga3_associate(normalized, ref)
    keys = jax_prng.split(key, pop_size + 1)                                                  # :165
    rbits = [jax_prng.randint_bits(k) for k in keys]
    sel = nsga3_niche(rank, last_rank, gid, gdist, len(ref), pop_size, rbits)
    return np.asarray(merged_pop)[sel], fit[sel], {"rank": rank, "last_rank": last_rank, "normalized": normalized,
                                                   "group_id": gid, "group_dist": gdist, "rbits": np.array(rbits, np.uint32),
                                                   "selected": sel}
