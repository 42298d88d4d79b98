"""NSGA-II variation kernel and the generation loop built from the hot-path operators."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _vary(pop, lb, ub, pro_c=1.0, dis_c=20.0, pro_m=1.0, dis_m=20.0, seed=1, gen=0):
    from evox_b200._lib import check, lib
    out = torch.empty_like(pop)
    check(lib().evx_variation_sbx_pm_f32(pop.data_ptr(), pop.shape[0], pop.shape[1], lb.data_ptr(), ub.data_ptr(), pro_c,
                                         dis_c, pro_m, dis_m, seed, gen, out.data_ptr(), None))
    torch.cuda.synchronize()
    return out


def test_sbx_algebra_without_mutation():
    """With pro_m = 0 the two children of a pair keep the parents' sum (sbx.py:27-34), stay parents
    when crossover is off (pro_c = 0 -> beta = 1, :23-25), and an odd last row passes through."""
    n, d = 20001, 12
    pop = torch.rand(n, d, device="cuda")
    lb, ub = torch.full((d,), -10.0, device="cuda"), torch.full((d,), 10.0, device="cuda")
    out = _vary(pop, lb, ub, pro_m=0.0)
    h = n // 2
    torch.testing.assert_close(out[:h] + out[h:2 * h], pop[:h] + pop[h:2 * h], rtol=1e-5, atol=1e-6)
    assert torch.equal(out[-1], pop[-1])
    same = _vary(pop, lb, ub, pro_c=0.0, pro_m=0.0)
    torch.testing.assert_close(same, pop, rtol=1e-6, atol=1e-7)
    # about half of the genes are copied unchanged (the beta3 < 0.5 -> 1 rule, sbx.py:22)
    frac = ((out[:h] - pop[:h]).abs() < 1e-7).float().mean().item()
    assert 0.45 < frac < 0.56


def test_sbx_beta_distribution_matches_reference_formula():
    """|beta| recovered from the children follows the reference's inverse-CDF transform of a uniform
    (sbx.py:16-18): compare empirical quantiles with a NumPy draw of the same formula."""
    n, d, dis_c = 200000, 4, 20.0
    pop = torch.rand(n, d, device="cuda")
    lb, ub = torch.full((d,), -100.0, device="cuda"), torch.full((d,), 100.0, device="cuda")
    out = _vary(pop, lb, ub, pro_m=0.0, dis_c=dis_c)
    h = n // 2
    num, den = (out[:h] - out[h:]), (pop[:h] - pop[h:])
    ok = den.abs() > 1e-3
    beta = (num[ok] / den[ok]).abs().cpu().numpy()
    beta = beta[np.abs(beta - 1.0) > 1e-4]                      # drop the "copied" half
    mu = np.random.default_rng(0).random(beta.size)
    ref = np.where(mu <= 0.5, (2 * mu) ** (1 / (dis_c + 1)), (2 - 2 * mu) ** (-1 / (dis_c + 1)))
    q = np.linspace(0.02, 0.98, 25)
    np.testing.assert_allclose(np.quantile(beta, q), np.quantile(ref, q), rtol=2e-2)


def test_polynomial_mutation_rate_and_bounds():
    n, d = 100000, 10
    pop = torch.rand(n, d, device="cuda") * 2 - 1
    lb, ub = torch.full((d,), -1.0, device="cuda"), torch.full((d,), 1.0, device="cuda")
    out = _vary(pop, lb, ub, pro_c=0.0, pro_m=1.0)              # crossover off: only mutation moves genes
    assert bool((out >= lb).all()) and bool((out <= ub).all())
    rate = ((out - pop).abs() > 1e-5).float().mean().item()        # beyond (p1+p2)/2 + (p1-p2)/2 rounding
    assert abs(rate - 1.0 / d) < 0.01                          # pm_mutation.py:16: site = U < pro_m / d
    a = _vary(pop, lb, ub, seed=7, gen=3)
    b = _vary(pop, lb, ub, seed=7, gen=3)
    c = _vary(pop, lb, ub, seed=7, gen=4)
    assert torch.equal(a, b) and not torch.equal(a, c)          # reproducible per (seed, generation)


@pytest.mark.parametrize("i", range(5))
def test_variation_values_match_reference_source(i):
    """Value parity for a given draw: the kernel's uniforms are regenerated on the CPU (oracle/philox.py),
    pushed through the reference's own `simulated_binary` / `polynomial` source on the jax shim
    (tests/golden/make_golden_variation.py, fixtures committed) and the kernel must reproduce the
    result.  Tolerance: 1e-6 relative plus 2e-6 of the gene's range - CUDA powf and NumPy's float32 pow
    differ by a few ulp, and pm_mutation.py:27-33 subtracts 1 from a power close to 1."""
    from pathlib import Path
    G = np.load(Path(__file__).parent / "golden" / "variation_outputs.npz")
    pc, dc, pm, dm = (float(v) for v in G[f"var{i}_par"])
    seed, gen = (int(v) for v in G[f"var{i}_seed"])
    pop, lb, ub = (torch.from_numpy(G[f"var{i}_{k}"]).cuda() for k in ("pop", "lb", "ub"))
    out = _vary(pop, lb, ub, pc, dc, pm, dm, seed, gen).cpu().numpy()
    want = G[f"var{i}_out"]
    span = (G[f"var{i}_ub"] - G[f"var{i}_lb"])[None, :]
    err = np.abs(out - want)
    assert (err <= 1e-6 * np.abs(want) + 2e-6 * span).all(), (i, float(err.max()))
    assert (out != G[f"var{i}_pop"]).any()                     # the case does exercise the operators


def test_variation_values_large_vs_oracle():
    """Same comparison at the configs[2] shape (pop 100k, d = 12) against the NumPy restatement
    (oracle_np.nsga2_variation, itself pinned to the reference source by tests/test_oracle_variation.py)."""
    from oracle import oracle_np as O
    n, d = 100_000, 12
    rng = np.random.default_rng(17)
    pop = rng.random((n, d), dtype=np.float32)
    lb, ub = np.zeros(d, np.float32), np.ones(d, np.float32)
    out = _vary(torch.from_numpy(pop).cuda(), torch.from_numpy(lb).cuda(), torch.from_numpy(ub).cuda(), seed=11, gen=250)
    want = O.nsga2_variation(pop, lb, ub, 1.0, 20.0, 1.0, 20.0, 11, 250)
    err = np.abs(out.cpu().numpy() - want)
    assert (err <= 1e-6 * np.abs(want) + 2e-6).all(), float(err.max())


def test_nsga2_converges_on_dtlz2():
    """configs[0]-style run (pop 100, M=3): after 100 generations the population sits on the unit
    sphere of DTLZ2 (sum f^2 = 1 on the true front) - an end-to-end sanity check of variation +
    evaluate + non_dominate, not a trajectory comparison (the reference asserts none either)."""
    import evox_b200 as E
    from evox_b200.algorithms import NSGA2, run_generations
    d, m = 12, 3
    alg = NSGA2(torch.zeros(d), torch.ones(d), n_objs=m, pop_size=100)
    prob = E.problems.numerical.DTLZ2(d=d, m=m)
    state = alg.setup(seed=3)
    r0 = None
    state, _ = run_generations(alg, prob, state, 1)
    r0 = (state.fitness ** 2).sum(1).sqrt().mean().item()
    state, _ = run_generations(alg, prob, state, 100)
    r = (state.fitness ** 2).sum(1).sqrt()
    assert state.population.shape == (100, d) and state.fitness.shape == (100, m)
    assert r.mean().item() < 1.08 < r0
    assert int(E.operators.non_dominated_sort(state.fitness).max()) <= 1
