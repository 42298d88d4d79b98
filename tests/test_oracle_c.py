"""The C restatement (oracle/oracle_c.c) against the NumPy oracle, the reference-run fixtures
and the reference's explicit known-answer vectors."""
import numpy as np

from oracle import oracle_c as C
from oracle import oracle_np as O


def test_c_nds_fixtures(golden):
    i = 0
    while f"nds{i}_x" in golden:
        assert (C.non_dominated_sort(golden[f"nds{i}_x"]) == golden[f"nds{i}_rank"]).all(), i
        i += 1
    x = np.array([[0.5, 4], [1, 2], [2, 1], [3, 1], [4, 1], [5, 0.5], [1, 4], [2, 3], [3, 2], [3, 3], [4, 4]],
                 np.float32)   # /root/reference/tests/test_non_dominated_sort.py:8-28
    assert C.non_dominated_sort(x).tolist() == [0, 0, 0, 1, 2, 0, 1, 1, 2, 3, 4]


def test_c_relation_fixtures(golden):
    i = 0
    while f"rel{i}_x" in golden:
        assert (C.dominate_relation(golden[f"rel{i}_x"], golden[f"rel{i}_y"]) == golden[f"rel{i}_d"]).all()
        i += 1


def test_c_crowding_fixtures(golden):
    i = 0
    while f"cd{i}_c" in golden:
        mask = golden[f"cd{i}_mask"] if bool(golden[f"cd{i}_has_mask"]) else None
        a, b = C.crowding_distance(golden[f"cd{i}_c"], mask), golden[f"cd{i}_d"]
        assert (np.isnan(a) == np.isnan(b)).all() and (a[np.isinf(b)] == b[np.isinf(b)]).all()
        f = np.isfinite(b)
        np.testing.assert_allclose(a[f], b[f], rtol=2e-6)
        i += 1


def test_c_vs_numpy_random():
    rng = np.random.default_rng(5)
    for n, m in [(700, 3), (500, 2), (400, 10), (1, 3), (0, 3)]:
        x = rng.random((n, m), dtype=np.float32)
        assert (C.non_dominated_sort(x) == O.non_dominated_sort(x)).all()
    x = rng.integers(0, 6, (600, 3)).astype(np.float32)
    assert (C.non_dominated_sort(x) == O.non_dominated_sort(x)).all()
    mask = rng.random(600) < 0.3
    a, b = C.crowding_distance(x, mask), O.crowding_distance(x, mask)
    assert np.array_equal(a, b, equal_nan=True)
