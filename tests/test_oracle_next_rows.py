"""Oracle pinning for the rows SURVEY.md section 8(f) marks "next" (no GPU, no /root/reference at run
time): the NumPy restatement against fixtures produced by executing the reference's own source on
the jax shim (tests/golden/make_golden_next.py).  f4 = SPEA2 fitness and truncation; the CUDA path for
this row does not exist yet - this is the checker it will be held to."""
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle_np as O

G = np.load(Path(__file__).parent / "golden" / "next_rows_outputs.npz")


@pytest.mark.parametrize("i", range(8))
def test_spea2_cal_fitness(i):
    obj, want = G[f"fit{i}_obj"], G[f"fit{i}_f"]
    got = O.spea2_cal_fitness(obj)
    assert got.dtype == np.float32 and got.shape == want.shape
    # raw fitness is an integer sum (exact); the density term is a float32 expression
    assert (np.floor(got) == np.floor(want)).all()
    np.testing.assert_allclose(got, want, rtol=1e-6, atol=0)


@pytest.mark.parametrize("i", range(5))
def test_spea2_truncation(i):
    obj, k, mask = G[f"trunc{i}_obj"], int(G[f"trunc{i}_k"]), G[f"trunc{i}_mask"]
    keep = O.spea2_truncation(obj, k, mask)
    assert (keep == G[f"trunc{i}_keep"]).all()
    assert int(mask.sum()) - int(keep.sum()) == k


def test_spea2_properties():
    rng = np.random.default_rng(3)
    obj = rng.random((200, 3), dtype=np.float32)
    f = O.spea2_cal_fitness(obj)
    nd = ~O.dominate_relation(obj, obj).any(0)                # non-dominated rows: raw fitness 0
    assert ((f < 1) == nd).all()                             # spea2.py:137 `mask = sig_fitness < 1`
    perm = rng.permutation(200)
    np.testing.assert_allclose(O.spea2_cal_fitness(obj[perm]), f[perm], rtol=1e-6)
