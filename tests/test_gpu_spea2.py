"""SPEA2 fitness and truncation through the C ABI, against the fixtures produced by
executing the reference's own spea2.py on the jax shim (tests/golden/make_golden_next.py) and against the
NumPy oracle at sizes the dense oracle still handles."""
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = np.load(Path(__file__).parent / "golden" / "next_rows_outputs.npz")


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("i", range(8))
def test_cal_fitness_fixtures(i):
    from evox_b200.algorithms.spea2 import cal_fitness
    obj, want = G[f"fit{i}_obj"], G[f"fit{i}_f"]
    got = cal_fitness(_cuda(obj)).cpu().numpy()
    assert got.dtype == np.float32
    assert (np.floor(got) == np.floor(want)).all()            # raw fitness is an integer sum: exact
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=0)  # the density term is a float32 expression


@pytest.mark.parametrize("i", range(5))
def test_truncation_fixtures(i):
    from evox_b200.algorithms.spea2 import _truncation
    obj, k, mask = G[f"trunc{i}_obj"], int(G[f"trunc{i}_k"]), G[f"trunc{i}_mask"]
    keep = _truncation(_cuda(obj), k, _cuda(mask)).cpu().numpy()
    assert (keep == G[f"trunc{i}_keep"]).all()


@pytest.mark.parametrize("n,m,kind", [(3000, 3, "random"), (2500, 2, "ties"), (1200, 10, "random"), (1, 3, "random"),
                                      (2, 3, "random"), (700, 5, "dupes")])
def test_cal_fitness_vs_oracle(n, m, kind):
    from oracle import oracle_np as O
    from evox_b200.algorithms.spea2 import cal_fitness
    rng = np.random.default_rng(n + m)
    obj = rng.random((n, m), dtype=np.float32)
    if kind == "ties":
        obj = rng.integers(0, 6, (n, m)).astype(np.float32)
    elif kind == "dupes":
        obj[n // 2:] = obj[:n - n // 2]
    want = O.spea2_cal_fitness(obj)
    got = cal_fitness(_cuda(obj)).cpu().numpy()
    assert (np.floor(got) == np.floor(want)).all()
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=0)


@pytest.mark.parametrize("n,m,k,p,kind", [(1500, 3, 700, 0.9, "random"), (800, 2, 300, 0.5, "ties"), (400, 3, 399, 1.0, "random"),
                                          (600, 4, 0, 0.7, "random"), (500, 3, 200, 1.0, "dupes"),
                                          # neighbour lists that run out (clusters of identical points, heavy removal)
                                          (900, 3, 600, 1.0, "clusters"), (1800, 3, 1000, 0.9, "random"),
                                          (700, 7, 650, 1.0, "random"), (600, 2, 400, 0.8, "clusters")])
def test_truncation_vs_oracle(n, m, k, p, kind):
    from oracle import oracle_np as O
    from evox_b200.algorithms.spea2 import _truncation
    rng = np.random.default_rng(n * 3 + k)
    obj = rng.random((n, m), dtype=np.float32)
    if kind == "ties":
        obj = rng.integers(0, 20, (n, m)).astype(np.float32)        # many equal distances: index order decides
    elif kind == "dupes":
        obj[::3] = obj[1::3][: len(obj[::3])]
    elif kind == "clusters":
        obj = obj[rng.integers(0, 20, n)]                            # 20 distinct points, ~n / 20 copies of each
    mask = rng.random(n) < p
    k = min(k, max(0, int(mask.sum()) - 1))
    want = O.spea2_truncation(obj, k, mask)
    got = _truncation(_cuda(obj), k, _cuda(mask)).cpu().numpy()
    assert (got == want).all(), np.nonzero(got != want)[0][:10]
    assert int(mask.sum()) - int(got.sum()) == k
