"""Oracle pinning for SURVEY.md section 8 row f2 (NSGA-II variation; no GPU, no /root/reference at run
time).  (i) The Philox4x32-10 restatement against the Random123 known-answer vectors.  (ii) The NumPy
restatement of sbx.py:8-40 + pm_mutation.py:7-54 + the clip of nsga2.py:85 against fixtures produced
by executing the reference's own source on the jax shim with the same uniforms
(tests/golden/make_golden_variation.py)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import oracle_np as O
from oracle import philox

G = np.load(Path(__file__).parent / "golden" / "variation_outputs.npz")
NCASES = 5


def test_philox_known_answers():
    # Random123 kat_vectors, philox4x32 with 10 rounds
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff, 0xffffffff), (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        got = philox.philox4x32_10(*ctr, *key)
        assert tuple(int(v) for v in got) == want


def test_draws_are_uniform_and_distinct():
    dr = philox.variation_draws(20000, 6, seed=3, generation=9)
    for k in ("sbx_mu", "sbx_keep", "pm_site", "pm_mu"):
        v = dr[k]
        assert v.dtype == np.float32 and 0.0 <= v.min() and v.max() < 1.0
        assert abs(float(v.mean()) - 0.5) < 0.01
    assert abs(float(dr["sbx_sign"].mean()) - 0.5) < 0.01
    assert abs(np.corrcoef(dr["sbx_mu"].ravel(), dr["sbx_keep"].ravel())[0, 1]) < 0.02


@pytest.mark.parametrize("i", range(NCASES))
def test_variation_oracle_matches_reference_source(i):
    pc, dc, pm, dm = (float(v) for v in G[f"var{i}_par"])
    seed, gen = (int(v) for v in G[f"var{i}_seed"])
    got = O.nsga2_variation(G[f"var{i}_pop"], G[f"var{i}_lb"], G[f"var{i}_ub"], pc, dc, pm, dm, seed, gen)
    want = G[f"var{i}_out"]
    assert got.dtype == np.float32 and got.shape == want.shape
    # both are NumPy float32 evaluations of the same expressions: equal to the last bit or two
    np.testing.assert_allclose(got, want, rtol=3e-7, atol=1e-7)
    assert (got >= G[f"var{i}_lb"]).all() and (got <= G[f"var{i}_ub"]).all()
