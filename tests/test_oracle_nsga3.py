"""Oracle pinning for SURVEY.md section 8 row f3 (NSGA-III environmental selection; no GPU, no
/root/reference at run time): the NumPy restatement of nsga3.py:97-210 against fixtures produced by
executing the reference's own `tell` on the jax shim (tests/golden/make_golden_nsga3.py).  The random
draws of the niching loop come from the Threefry restatement in oracle/jax_prng.py on both sides: the
fixtures pin the selection logic, not jaxlib's `split` / `randint` (no reference test holds a vector for
those: that part stays unpinned)."""
from pathlib import Path

import numpy as np
import pytest

from oracle import jax_prng
from oracle import oracle_np as O

G = np.load(Path(__file__).parent / "golden" / "nsga3_outputs.npz")


@pytest.mark.parametrize("i", range(5))
def test_nsga3_tell_matches_reference_source(i):
    pop, fit, ref = G[f"n3_{i}_pop"], G[f"n3_{i}_fit"], G[f"n3_{i}_ref"]
    key = tuple(np.uint32(v) for v in G[f"n3_{i}_key"])
    p, f, aux = O.nsga3_tell(pop, fit, ref, int(G[f"n3_{i}_popsize"]), key)
    assert (p == G[f"n3_{i}_opop"]).all() and (f == G[f"n3_{i}_ofit"]).all()
    assert aux["selected"].shape[0] == int(G[f"n3_{i}_popsize"])


def test_randint_arithmetic():
    # span 1 always gives minval; results stay inside [minval, maxval)
    assert jax_prng.randint_from_bits(0xFFFFFFFF, 0xFFFFFFFF, 0, 1) == 0
    rng = np.random.default_rng(0)
    for span in (2, 3, 7, 1000, 65537, 2**31 - 1):
        for _ in range(50):
            hi, lo = (int(v) for v in rng.integers(0, 2**32, 2))
            r = jax_prng.randint_from_bits(hi, lo, 0, span)
            assert 0 <= r < span
            # the published formula: ((hi % span) * ((2^16 % span)^2 % span) + lo % span) % span in uint32
            mult = (pow(2, 16, span) ** 2) % span
            assert r == ((((hi % span) * mult) & 0xFFFFFFFF) + lo % span) % span or span > 65536
    keys = jax_prng.split(jax_prng.prng_key(7), 5)
    assert len(keys) == 5 and len({(int(a), int(b)) for a, b in keys}) == 5
