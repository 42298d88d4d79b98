"""Models (no GPU) of the two sequential loops the CUDA kernels restructure, replayed in NumPy and held to the
oracle's step-by-step restatements:

* NSGA-III niching (`csrc/nsga3.cu`, nsga3.py:156-201): the reference serves, one pick per iteration, the first
  niche with the smallest count.  The kernel serves all niches at the smallest count in one step, in index
  order; a pick's loop index - which selects its random words - is the number selected so far plus the niche's
  position in the level.
* SPEA2 truncation (`csrc/spea2.cu`, spea2.py:40-67): the reference recomputes every row's distance to its
  nearest remaining neighbour after every removal.  The kernel keeps, per row, a sorted list of its eight
  nearest neighbours and repairs a row only when it reaches the arg-min with a removed neighbour on show.
"""
import numpy as np
import pytest

from oracle import jax_prng
from oracle import oracle_np as O

F32 = np.float32


def niche_by_levels(rank, last_rank, group_id, group_dist, nref, pop_size, rbits):
    rank = np.array(rank, np.int64, copy=True)
    last = rank == last_rank
    rho = np.bincount(np.where(rank < last_rank, group_id, nref), minlength=nref + 1)[:nref].astype(np.int64)
    members = [np.nonzero(last & (group_id == g))[0] for g in range(nref)]            # index order
    left = np.array([len(m) for m in members])
    taken = [np.zeros(len(m), bool) for m in members]
    INF = np.iinfo(np.int64).max
    rho = np.where(left == 0, INF, rho)
    selected = int((rank < last_rank).sum())
    while selected < pop_size:
        level = rho.min()
        if level == INF:
            break
        active = np.nonzero(rho == level)[0]                                          # ascending niche index
        for pos, g in enumerate(active):
            loop_index = selected + pos
            if loop_index >= pop_size:
                break
            remaining = np.nonzero(~taken[g])[0]
            if level == 0 and left[g] > 1:
                d = group_dist[members[g][remaining]]
                k = int(np.argmin(d))                                                 # first minimum: lowest index
            elif left[g] == 1:
                k = 0
            else:
                hi, lo = rbits[loop_index]
                k = jax_prng.randint_from_bits(hi, lo, 0, int(left[g]))
            q = remaining[k]
            taken[g][q] = True
            rank[members[g][q]] = last_rank - 1
            left[g] -= 1
            rho[g] = INF if left[g] == 0 else level + 1
        selected += min(len(active), pop_size - selected)
    return np.nonzero(rank < last_rank)[0][:pop_size].astype(np.int32)


@pytest.mark.parametrize("n,nref,ps,seed", [(400, 30, 200, 0), (500, 5, 250, 1), (300, 2, 100, 2), (600, 90, 300, 3),
                                            (350, 1, 120, 4), (450, 40, 60, 5)])
def test_niche_by_levels_equals_sequential_loop(n, nref, ps, seed):
    rng = np.random.default_rng(seed)
    rank = np.sort(rng.integers(0, 6, n)).astype(np.int32)[rng.permutation(n)]
    last_rank = int(rank[np.argsort(rank, kind="stable")[ps]])
    gid = np.minimum(rng.geometric(0.2, n) - 1, nref - 1).astype(np.int32)            # crowded and thin niches
    gid[rank > last_rank] = nref
    gd = (rng.integers(0, 20, n) / 20).astype(F32)                                     # ties in distance
    gd[rank > last_rank] = np.nan
    rbits = [(int(a), int(b)) for a, b in rng.integers(0, 2**32, (ps + 1, 2), dtype=np.uint64)]
    want = O.nsga3_niche(rank, last_rank, gid, gd, nref, ps, rbits)
    got = niche_by_levels(rank, last_rank, gid, gd, nref, ps, rbits)
    assert (got == want).all()


def truncation_lazy(obj, k, mask, knn=8):
    obj = np.asarray(obj, F32)
    mask = np.asarray(mask, bool)
    n = obj.shape[0]
    dis = O.pairwise_euclidean_dist(obj, obj).astype(F32)
    dis[np.arange(n), np.arange(n)] = np.inf
    dis = np.where(mask[:, None] & mask[None, :], dis, F32(np.inf)).astype(F32)
    alive = mask.copy()
    lists = []
    for i in range(n):
        cand = np.nonzero(mask & (np.arange(n) != i))[0] if mask[i] else np.array([], int)
        order = cand[np.lexsort((cand, dis[i, cand]))][:knn]                           # by (distance, index)
        lists.append(list(order))
    ptr = np.zeros(n, int)
    nn = np.array([dis[i, lists[i][0]] if lists[i] else np.inf for i in range(n)], F32)
    nni = np.array([lists[i][0] if lists[i] else -1 for i in range(n)])
    removed = 0
    while removed < k:
        shown = np.where(alive, nn, F32(np.inf))
        idx = int(np.argsort(shown, kind="stable")[0])                                # ties: lowest index
        if alive[idx] and np.isfinite(nn[idx]) and not alive[nni[idx]]:                # neighbour on show is gone
            p = ptr[idx] + 1
            while p < len(lists[idx]) and not alive[lists[idx][p]]:
                p += 1
            if p < len(lists[idx]):
                ptr[idx] = p
            else:                                                                      # list used up: search again
                cand = np.nonzero(alive & (np.arange(n) != idx))[0]
                if len(cand):
                    best = cand[np.lexsort((cand, dis[idx, cand]))][0]
                    lists[idx], ptr[idx] = [best], 0
                else:
                    lists[idx], ptr[idx] = [], 0
            nn[idx] = dis[idx, lists[idx][ptr[idx]]] if lists[idx] else np.inf
            nni[idx] = lists[idx][ptr[idx]] if lists[idx] else -1
            continue
        partner = nni[idx] if (alive[idx] and np.isfinite(nn[idx])) else -1
        alive[idx] = False
        nn[idx] = np.inf
        removed += 1
        # the removed row's nearest neighbour, if the two were each other's nearest, is moved on in the same turn
        # (a used-up list is left to the lazy repair above)
        if partner >= 0 and alive[partner] and nni[partner] == idx:
            p = ptr[partner] + 1
            while p < len(lists[partner]) and not alive[lists[partner][p]]:
                p += 1
            if p < len(lists[partner]):
                ptr[partner] = p
                nn[partner] = dis[partner, lists[partner][p]]
                nni[partner] = lists[partner][p]
    return alive


@pytest.mark.parametrize("n,m,k,p,kind", [(120, 3, 60, 1.0, "random"), (150, 2, 100, 0.8, "ties"), (90, 3, 80, 1.0, "clusters"),
                                          (100, 4, 0, 0.7, "random"), (80, 3, 30, 0.5, "dupes")])
def test_truncation_lazy_lists_equal_full_recomputation(n, m, k, p, kind):
    rng = np.random.default_rng(n + k)
    obj = rng.random((n, m), dtype=F32)
    if kind == "ties":
        obj = rng.integers(0, 6, (n, m)).astype(F32)
    elif kind == "clusters":
        obj = obj[rng.integers(0, 6, n)]
    elif kind == "dupes":
        obj[::3] = obj[1::3][: len(obj[::3])]
    mask = rng.random(n) < p
    k = min(k, max(0, int(mask.sum()) - 1))
    want = O.spea2_truncation(obj, k, mask)
    got = truncation_lazy(obj, k, mask)
    assert (got == want).all()
