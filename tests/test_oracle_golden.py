"""Oracle (oracle/oracle_np.py) against fixtures produced by executing the reference's own
source files on the jax shim (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle import oracle_np as O

RTOL = 2e-6   # float32 re-association only; integer results are compared exactly


def _cases(g, prefix, field):
    i = 0
    while f"{prefix}{i}_{field}" in g:
        yield i
        i += 1


def test_nds_matches_reference(golden):
    n = 0
    for i in _cases(golden, "nds", "x"):
        x = golden[f"nds{i}_x"]
        assert (O.non_dominated_sort(x) == golden[f"nds{i}_rank"]).all(), i
        n += 1
    assert n >= 19


def test_dominate_relation_matches_reference(golden):
    for i in _cases(golden, "rel", "x"):
        d = O.dominate_relation(golden[f"rel{i}_x"], golden[f"rel{i}_y"])
        assert d.dtype == bool and (d == golden[f"rel{i}_d"]).all(), i


def _close(a, b, rtol=RTOL):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    assert (np.isnan(a) == np.isnan(b)).all()
    inf = np.isinf(b)
    assert (a[inf] == b[inf]).all()
    ok = np.isfinite(b)
    np.testing.assert_allclose(a[ok], b[ok], rtol=rtol, atol=0)


def test_crowding_matches_reference(golden):
    n = 0
    for i in _cases(golden, "cd", "c"):
        c = golden[f"cd{i}_c"]
        mask = golden[f"cd{i}_mask"] if bool(golden[f"cd{i}_has_mask"]) else None
        _close(O.crowding_distance(c, mask), golden[f"cd{i}_d"])
        assert (O.crowding_distance_sort(c, mask) == golden[f"cd{i}_sort"]).all(), i
        n += 1
    assert n >= 17


def test_non_dominate_matches_reference(golden):
    for i in _cases(golden, "sel", "pop"):
        p, f = O.non_dominate(golden[f"sel{i}_pop"], golden[f"sel{i}_fit"], int(golden[f"sel{i}_topk"]),
                              bool(golden[f"sel{i}_dedup"]))
        assert (p == golden[f"sel{i}_opop"]).all(), i
        assert (f == golden[f"sel{i}_ofit"]).all(), i


@pytest.mark.parametrize("k", range(1, 8))
def test_dtlz_matches_reference(golden, k):
    n = 0
    for j in range(5):
        if f"dtlz{k}_{j}_X" not in golden:
            continue
        F = O.dtlz(k, golden[f"dtlz{k}_{j}_X"], int(golden[f"dtlz{k}_{j}_m"]))
        _close(F, golden[f"dtlz{k}_{j}_F"], rtol=1e-5)
        n += 1
    assert n >= 4


def test_maf1_matches_reference(golden):
    for j in range(3):
        _close(O.maf1(golden[f"maf1_{j}_X"], int(golden[f"maf1_{j}_m"])), golden[f"maf1_{j}_F"])


@pytest.mark.parametrize("k", range(1, 10))
def test_lsmop_matches_reference(golden, k):
    for j in range(4):
        X, m = golden[f"lsmop{k}_{j}_X"], int(golden[f"lsmop{k}_{j}_m"])
        sub, off = O.lsmop_groups(X.shape[1], m)
        assert list(sub) == golden[f"lsmop{k}_{j}_sublen"].tolist()
        assert list(off) == golden[f"lsmop{k}_{j}_off"].tolist()
        _close(O.lsmop(k, X, m), golden[f"lsmop{k}_{j}_F"], rtol=1e-5)


def test_nds_properties():
    rng = np.random.default_rng(3)
    x = rng.random((300, 3), dtype=np.float32)
    r = O.non_dominated_sort(x)
    perm = rng.permutation(300)
    assert (O.non_dominated_sort(x[perm]) == r[perm]).all()          # permutation invariance
    d = O.dominate_relation(x, x)
    i, j = np.nonzero(d)
    assert (r[i] < r[j]).all()                                       # a dominates b => rank(a) < rank(b)
    x2 = np.concatenate([x, x[:50]])
    r2 = O.non_dominated_sort(x2)
    assert (r2[300:] == r2[:50]).all() and (r2[:300] == r).all()     # duplicates share a rank
