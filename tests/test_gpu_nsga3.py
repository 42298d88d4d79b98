"""NSGA-III association and niching through the C ABI against the NumPy oracle
(itself pinned to the reference's `tell` by tests/test_oracle_nsga3.py) and, end to end, against the
fixtures produced by executing the reference's own source on the jax shim."""
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
G = np.load(Path(__file__).parent / "golden" / "nsga3_outputs.npz")


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _assert_dist_close(x, got, want):
    """d = |x| * sqrt(1 - c^2) loses digits as c -> 1 (a float32 rounding of c moves 1 - c^2 by 1.2e-7
    whatever its size), so the distances are compared where the formula is well conditioned: on
    (d / |x|)^2 = 1 - c^2, to 1.5e-6 absolute (an m-term float32 dot product summed in another order - NumPy
    hands it to BLAS - plus the two normalisations)."""
    nx = np.linalg.norm(x.astype(np.float64), axis=1)
    a, b = (got / nx) ** 2, (want / nx) ** 2
    assert np.abs(a - b).max() <= 1.5e-6, float(np.abs(a - b).max())


def _case(i):
    from oracle import oracle_np as O
    pop, fit, ref = G[f"n3_{i}_pop"], G[f"n3_{i}_fit"], G[f"n3_{i}_ref"]
    key = tuple(np.uint32(v) for v in G[f"n3_{i}_key"])
    ps = int(G[f"n3_{i}_popsize"])
    return pop, fit, ref, ps, O.nsga3_tell(pop, fit, ref, ps, key)


@pytest.mark.parametrize("i", range(5))
def test_associate_and_niche_on_fixture_cases(i):
    """Association within float32 rounding of the cosine (see _assert_dist_close); the niching
    loop, fed the oracle's association, must return exactly the reference's survivors."""
    from evox_b200.algorithms import nsga3
    pop, fit, ref, ps, (opop, ofit, aux) = _case(i)
    gid, gd = nsga3.associate(_cuda(aux["normalized"]), _cuda(ref))
    gid, gd = gid.cpu().numpy(), gd.cpu().numpy()
    want_d = aux["group_dist"]
    assert (np.isnan(gd) == np.isnan(want_d)).all()
    ok = ~np.isnan(want_d)
    _assert_dist_close(aux["normalized"][ok], gd[ok], want_d[ok])
    assert (gid[~ok] == len(ref)).all()
    assert (gid[ok] == aux["group_id"][ok]).mean() > 0.97               # only near-ties between two directions may differ
    sel = nsga3.niche(_cuda(aux["rank"].astype(np.int32)), aux["last_rank"], _cuda(aux["group_id"]), _cuda(aux["group_dist"]),
                      len(ref), ps, _cuda(aux["rbits"].view(np.int32)))
    assert (sel.cpu().numpy() == aux["selected"]).all()
    assert (pop[aux["selected"]] == G[f"n3_{i}_opop"]).all() and (fit[aux["selected"]] == G[f"n3_{i}_ofit"]).all()


@pytest.mark.parametrize("n,m,nref,ps", [(4000, 3, 91, 2000), (3000, 5, 210, 1500), (2000, 10, 275, 1000), (1000, 2, 1000, 500),
                                         # niches with more than 64 candidates (served by the whole CTA): random picks,
                                         # then (last front = front 0, every count 0) closest-candidate picks as well
                                         (4000, 3, 5, 2000), (3000, 3, 2, 1500), (2000, 3, 3, 200), (6000, 4, 1, 2500)])
def test_niche_larger_vs_oracle(n, m, nref, ps):
    """Random but consistent inputs at sizes of a few thousand: ranks with a last front that must be cut,
    random group assignments and distances (with ties), random draw words."""
    from oracle import oracle_np as O
    from evox_b200.algorithms import nsga3
    rng = np.random.default_rng(n + m + nref)
    rank = np.sort(rng.integers(0, 6, n)).astype(np.int32)[rng.permutation(n)]
    order = np.argsort(rank, kind="stable")
    last_rank = int(rank[order[ps]])
    gid = rng.integers(0, nref, n).astype(np.int32)
    gid[rank > last_rank] = nref
    gd = (rng.integers(0, 50, n) / 50).astype(np.float32)
    gd[rank > last_rank] = np.nan
    rbits = rng.integers(0, 2**32, (ps + 1, 2), dtype=np.uint64).astype(np.uint32)
    want = O.nsga3_niche(rank, last_rank, gid, gd, nref, ps, [(int(a), int(b)) for a, b in rbits])
    got = nsga3.niche(_cuda(rank), last_rank, _cuda(gid), _cuda(gd), nref, ps, _cuda(rbits.view(np.int32)))
    assert (got.cpu().numpy() == want).all()


def test_associate_large_vs_oracle():
    from oracle import oracle_np as O
    from evox_b200.algorithms import nsga3
    rng = np.random.default_rng(5)
    x = rng.random((20000, 10), dtype=np.float32)
    x[::7] = np.nan
    ref = rng.random((500, 10), dtype=np.float32) + np.float32(0.01)
    gid, gd = nsga3.associate(_cuda(x), _cuda(ref))
    wid, wd = O.nsga3_associate(x, ref)
    gid, gd = gid.cpu().numpy(), gd.cpu().numpy()
    ok = ~np.isnan(wd)
    assert (np.isnan(gd) == ~ok).all() and (gid[~ok] == 500).all()
    _assert_dist_close(x[ok], gd[ok], wd[ok])
    assert (gid[ok] == wid[ok]).mean() > 0.995


@pytest.mark.parametrize("i", range(5))
def test_environmental_selection_end_to_end(i):
    """The whole NSGA3.tell mirror (library sort -> torch normalisation -> association -> niching).  The
    normalisation's M x M solve runs in another library than the reference's, so survivors are compared as
    sets with a small allowance for rows whose nearest direction is a near-tie."""
    from evox_b200.algorithms import nsga3
    pop, fit, ref, ps, (opop, ofit, aux) = _case(i)
    p, f = nsga3.environmental_selection(_cuda(pop), _cuda(fit), _cuda(ref), ps, _cuda(aux["rbits"].view(np.int32)))
    assert p.shape == (ps, pop.shape[1]) and f.shape == (ps, fit.shape[1])
    got = {tuple(r) for r in f.cpu().numpy().tolist()}
    want = {tuple(r) for r in ofit.tolist()}
    assert len(got & want) >= len(want) - max(2, ps // 10)


def test_niche_repeatable():
    """Same reason as test_nds_small_population_repeatable: the niching loop is one CTA with hand-placed barriers
    and warps that serve heavy niches side by side."""
    from evox_b200.algorithms import nsga3
    rng = np.random.default_rng(11)
    n, nref, ps = 6000, 40, 3000
    rank = np.sort(rng.integers(0, 6, n)).astype(np.int32)[rng.permutation(n)]
    last_rank = int(rank[np.argsort(rank, kind="stable")[ps]])
    gid = np.minimum(rng.geometric(0.15, n) - 1, nref - 1).astype(np.int32)      # a few crowded niches, many thin ones
    gid[rank > last_rank] = nref
    gd = (rng.integers(0, 50, n) / 50).astype(np.float32)
    gd[rank > last_rank] = np.nan
    rbits = rng.integers(0, 2**32, (ps + 1, 2), dtype=np.uint64).astype(np.uint32)
    from oracle import oracle_c  # noqa: F401  (build check only)
    from oracle import oracle_np as O
    want = O.nsga3_niche(rank, last_rank, gid, gd, nref, ps, [(int(a), int(b)) for a, b in rbits])
    args = (_cuda(rank), last_rank, _cuda(gid), _cuda(gd), nref, ps, _cuda(rbits.view(np.int32)))
    for _ in range(25):
        assert (nsga3.niche(*args).cpu().numpy() == want).all()
