"""Model of the fused survivor selection planned for SURVEY.md section 8 row f1 (no GPU): the
decomposition a CUDA version will use - rank histogram -> cut front -> crowding distance on the cut
front ALONE -> permutation by compaction + one small sort -> one key sort of the survivors - replayed
in NumPy and held to the oracle's restatement of non_dominate.py:210-222, including the :216/:218 quirk
(the secondary key of row i is entry i of the crowding argsort PERMUTATION, not row i's distance)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle_np as O


def select_model(fit, rank, topk):
    n = rank.shape[0]
    hist = np.bincount(rank, minlength=n)                          # ranks are < n
    worst = int(np.searchsorted(np.cumsum(hist), topk, side="left"))   # first rank whose cumulative count reaches topk
    cut = rank == worst
    cut_idx = np.nonzero(cut)[0]                                   # stream compaction, index order
    rest_idx = np.nonzero(~cut)[0]
    d = O.crowding_distance(fit[cut_idx])                          # the cut front by itself, no mask
    # stable ascending argsort with NaN last == jnp.argsort; rows outside the cut front carry -inf and come first
    by_d = cut_idx[np.argsort(d, kind="stable")]
    cd_perm = np.concatenate([rest_idx, by_d]).astype(np.int64)
    key = rank.astype(np.int64) * n + (n - 1 - cd_perm)           # rank ascending, then cd_perm[i] descending
    key[rank > worst] = np.iinfo(np.int64).max
    return np.argsort(key, kind="stable")[:topk], worst, cd_perm


def oracle_combined(fit, rank, topk):
    order = np.argsort(rank, kind="stable")
    worst = rank[order[topk - 1]]
    cd_perm = O.crowding_distance_sort(fit, rank == worst)
    return np.lexsort((-cd_perm, rank))[:topk], int(worst), cd_perm


@pytest.mark.parametrize("n,m,topk,kind", [(200, 3, 100, "random"), (200, 3, 1, "random"), (200, 3, 200, "random"),
                                           (301, 2, 150, "random"), (120, 5, 60, "random"), (150, 3, 75, "ties"),
                                           (90, 3, 45, "one_front"), (60, 2, 30, "chain"), (80, 3, 40, "flat_objective")])
def test_model_equals_oracle(n, m, topk, kind):
    rng = np.random.default_rng(n * 7 + m + topk)
    fit = rng.random((n, m), dtype=np.float32)
    if kind == "ties":
        fit = rng.integers(0, 4, (n, m)).astype(np.float32)
    elif kind == "one_front":
        fit = (fit / fit.sum(1, keepdims=True)).astype(np.float32)
    elif kind == "chain":
        fit = (np.arange(n, dtype=np.float32)[:, None] * np.ones((1, m), np.float32))[rng.permutation(n)]
    elif kind == "flat_objective":
        fit[:, 1] = 0.5                                            # zero range: 0/0 = NaN distances, sorted last
    rank = O.non_dominated_sort(fit)
    got, w1, p1 = select_model(fit, rank, topk)
    want, w2, p2 = oracle_combined(fit, rank, topk)
    assert w1 == w2
    assert (p1 == p2).all()
    assert (got == want).all()


def test_model_on_reference_fixtures():
    G = np.load(Path(__file__).parent / "golden" / "reference_outputs.npz")
    for i in range(6):
        if bool(G[f"sel{i}_dedup"]):
            continue
        pop, fit, topk = G[f"sel{i}_pop"], G[f"sel{i}_fit"], int(G[f"sel{i}_topk"])
        rank = O.non_dominated_sort(fit)
        idx, _, _ = select_model(fit, rank, topk)
        assert (pop[idx] == G[f"sel{i}_opop"]).all() and (fit[idx] == G[f"sel{i}_ofit"]).all()
