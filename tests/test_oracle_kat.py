"""The reference's own known-answer tests for this path, run against the oracle.

Inputs with explicit values are restated from the cited reference tests; inputs drawn
from jax.random are regenerated bit-for-bit by oracle/jax_prng.py (Threefry-2x32)."""
import re

import numpy as np
import pytest

from oracle import jax_prng
from oracle import oracle_np as O

INF = np.inf


def test_nds_reference_kat():
    # /root/reference/tests/test_non_dominated_sort.py:8-28
    x = np.array([[0.5, 4], [1, 2], [2, 1], [3, 1], [4, 1], [5, 0.5], [1, 4], [2, 3], [3, 2],
                  [3, 3], [4, 4]], np.float32)
    assert O.non_dominated_sort(x).tolist() == [0, 0, 0, 1, 2, 0, 1, 1, 2, 3, 4]
    assert O.non_dominated_sort(x).dtype == np.int32


def test_crowding_reference_kat():
    # /root/reference/tests/test_crowding_distance.py:7-21
    x = np.array([[0.5, 4], [1, 2.5], [2, 2], [3, 1], [4, 0.8]], np.float32)
    exp = np.array([INF, 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, 2 / 3.5 + 1.2 / 3.2, INF], np.float32)
    np.testing.assert_allclose(O.crowding_distance(x), exp, rtol=1e-6)


def test_masked_crowding_reference_kat():
    # /root/reference/tests/test_crowding_distance.py:127-146
    x = np.array([[-1, -1], [0.5, 4], [1, 2.5], [2, 2], [-2, -2], [3, 1], [4, 0.8], [-3, -3], [-3, -4]],
                 np.float32)
    mask = np.array([False, True, True, True, False, True, True, False, False])
    exp = np.array([-INF, INF, 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, -INF, 2 / 3.5 + 1.2 / 3.2, INF,
                    -INF, -INF], np.float32)
    np.testing.assert_allclose(O.crowding_distance(x, mask), exp, rtol=1e-6)


def test_crowding_prng_golden(golden):
    # /root/reference/tests/test_crowding_distance.py:24-124: normal(PRNGKey(314),(128,8)) ->
    # rank-0 subset -> crowding distance, rtol=1e-5.  The 90 expected values of :28-121 are
    # restated in tests/golden/kat_crowding2.npy.
    x = jax_prng.normal(jax_prng.prng_key(314), (128, 8))
    pf = x[O.non_dominated_sort(x) == 0]
    d = O.crowding_distance(pf)
    exp = _kat_cd2()
    assert d.shape == exp.shape == (90,)
    assert (np.isinf(d) == np.isinf(exp)).all()
    f = np.isfinite(exp)
    np.testing.assert_allclose(d[f], exp[f], rtol=1e-5)


def _kat_cd2():
    from pathlib import Path
    return np.load(Path(__file__).parent / "golden" / "kat_crowding2.npy")


def test_monitor_pareto_archive_kat():
    # /root/reference/tests/test_monitors.py:30-48 (StdMOMonitor keeps rank==0 of the archive,
    # src/evox/monitors/std_mo_monitor.py:89); integer-valued fitness.
    f1 = np.array([[1, 2], [3, 1], [5, 6]])
    keep = O.non_dominated_sort(f1) == 0
    assert (f1[keep] == np.array([[1, 2], [3, 1]])).all()
    f2 = np.concatenate([f1[keep], np.array([[0.5, 1.5], [7, 8], [0, 10]])])
    keep2 = O.non_dominated_sort(f2) == 0
    assert (f2[keep2] == np.array([[3, 1], [0.5, 1.5], [0, 10]])).all()


def test_dtlz1_reference_kat():
    # /root/reference/tests/test_classic_problems.py:46-54
    F = O.dtlz(1, np.ones((16, 7), np.float32) * 0.5, 4)
    np.testing.assert_allclose(F.sum(axis=1), 0.5 * np.ones(16), atol=1e-6)


_LSMOP_GOLD = {   # /root/reference/tests/test_lsmop.py:19-124: (r1[1,1], r1[1,2], relative?)
    1: (12.5981, 21.2454, False, False), 2: (0.6876, 0.4688, False, False),
    3: (5.3782496e04, 26.0379, True, False), 4: (0.5550, 0.8865, False, False),
    5: (2.6036, 10.7802, False, False), 6: (1.7095365e03, 2.5482053e04, True, True),
    7: (1.4722900e04, 1.3406048, True, False), 8: (1.9358, 0.8251, False, False),
    9: (0.0523, 217.5438, False, False),
}


@pytest.mark.parametrize("k", range(1, 10))
def test_lsmop_prng_golden(k):
    # /root/reference/tests/test_lsmop.py:10-14: uniform(PRNGKey(0),(100,300),0,[1,1,10,...])
    upper = np.concatenate([np.ones(2, np.float32), 10 * np.ones(298, np.float32)])
    data = jax_prng.uniform(jax_prng.prng_key(0), (100, 300), 0.0, upper)
    r = O.lsmop(k, data, 3)
    assert r.shape == (100, 3)
    a, b, rel_a, rel_b = _LSMOP_GOLD[k]
    for got, exp, rel in ((r[1, 1], a, rel_a), (r[1, 2], b, rel_b)):
        if rel:
            assert abs(float(got) - exp) / exp < 1e-5
        else:
            assert abs(float(got) - exp) < 1e-4


def test_maf1_prng_golden():
    # /root/reference/tests/test_maf.py:10-11, :49-57
    data = jax_prng.uniform(jax_prng.prng_key(1), (3, 12))
    r = O.maf1(data, 3)
    assert r.shape == (3, 3)
    assert abs(float(r[1, 1]) - 1.8404) < 1e-4


def test_threefry_known_answer():
    # Random123 / jax tests/random_test.py known-answer vectors for Threefry-2x32-20.
    x0, x1 = jax_prng.threefry2x32(0x0, 0x0, np.array([0], np.uint32), np.array([0], np.uint32))
    assert (int(x0[0]), int(x1[0])) == (0x6B200159, 0x99BA4EFE)
    x0, x1 = jax_prng.threefry2x32(0xFFFFFFFF, 0xFFFFFFFF, np.array([0xFFFFFFFF], np.uint32),
                                   np.array([0xFFFFFFFF], np.uint32))
    assert (int(x0[0]), int(x1[0])) == (0x1CB996FC, 0xBB002BE7)
    x0, x1 = jax_prng.threefry2x32(0x13198A2E, 0x03707344, np.array([0x243F6A88], np.uint32),
                                   np.array([0x85A308D3], np.uint32))
    assert (int(x0[0]), int(x1[0])) == (0xC4923A9C, 0x483DF7A0)
