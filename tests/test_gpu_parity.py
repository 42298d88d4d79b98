"""Parity of the CUDA path (through the C ABI, via the evox_b200 host mirror) against the oracle, the
reference-run fixtures and the reference's known-answer vectors.  Ranks / masks / selections are
compared bit-exactly; crowding distances to 1e-6 relative with exact infinities (north star);
problem values to 2e-5 relative (float32 re-association and libm-vs-XLA transcendental ulps)."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

CD_RTOL = 1e-6
EVAL_RTOL = 2e-5


@pytest.fixture(scope="module")
def E():
    import evox_b200
    evox_b200.lib()
    return evox_b200


def _cuda(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _close(a, b, rtol, atol=0.0):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape
    assert (np.isnan(a) == np.isnan(b)).all(), "NaN pattern differs"
    inf = np.isinf(b)
    assert (a[inf] == b[inf]).all(), "infinities differ"
    ok = np.isfinite(b)
    np.testing.assert_allclose(a[ok], b[ok], rtol=rtol, atol=atol)


# ------------------------------------------------------------------------------------------------
# non_dominated_sort
# ------------------------------------------------------------------------------------------------
def test_nds_reference_kat(E):
    # /root/reference/tests/test_non_dominated_sort.py:8-28
    x = np.array([[0.5, 4], [1, 2], [2, 1], [3, 1], [4, 1], [5, 0.5], [1, 4], [2, 3], [3, 2], [3, 3], [4, 4]],
                 np.float32)
    for method in ("auto", "scan", "full map-reduce", "host"):
        r = E.operators.non_dominated_sort(_cuda(x), method)
        assert r.dtype == torch.int32
        assert r.cpu().tolist() == [0, 0, 0, 1, 2, 0, 1, 1, 2, 3, 4]


def test_nds_fixtures(E, golden):
    i = 0
    while f"nds{i}_x" in golden:
        x = golden[f"nds{i}_x"]
        r = E.operators.non_dominated_sort(_cuda(x)).cpu().numpy()
        assert (r == golden[f"nds{i}_rank"]).all(), (i, x.shape, r[:20], golden[f"nds{i}_rank"][:20])
        i += 1
    assert i >= 19


def test_nds_monitor_integer_inputs(E):
    # /root/reference/tests/test_monitors.py:30-48: integer fitness through the same operator
    f1 = torch.tensor([[1, 2], [3, 1], [5, 6]])
    assert (E.operators.non_dominated_sort(f1).cpu().numpy() == [0, 0, 1]).all()


@pytest.mark.parametrize("n,m", [(1, 3), (2, 3), (31, 2), (1023, 3), (1024, 3), (1025, 3), (2049, 2), (3000, 5),
                                 (5000, 3), (4097, 10), (20000, 3), (6000, 1), (2500, 13), (1500, 32)])
def test_nds_random_vs_oracle(E, n, m):
    from oracle import oracle_c as C
    rng = np.random.default_rng(n * 31 + m)
    x = rng.random((n, m), dtype=np.float32)
    r = E.operators.non_dominated_sort(_cuda(x)).cpu().numpy()
    ref = C.non_dominated_sort(x)
    bad = np.nonzero(r != ref)[0]
    assert bad.size == 0, (n, m, bad[:10], r[bad[:10]], ref[bad[:10]])


@pytest.mark.parametrize("n,m,kind", [(8192, 3, "random"), (8193, 3, "random"), (7000, 6, "random"), (7000, 7, "random"),
                                      (8000, 3, "ties"), (8192, 2, "one_front"), (6000, 3, "nan")])
def test_nds_small_population_paths_agree(E, n, m, kind, monkeypatch):
    """Up to 8192 rows the sort takes two single-CTA kernels (k_small_sort_unique while the keys fit in shared memory - M <= 6 at 8192 rows -, k_peel_single);
    EVX_PREP=radix / EVX_PEEL=dynamic switch them off.  All combinations and the oracle must agree, at the
    limits of both kernels."""
    from oracle import oracle_c as C
    rng = np.random.default_rng(n + m)
    x = rng.random((n, m), dtype=np.float32)
    if kind == "ties":
        x = rng.integers(0, 15, (n, m)).astype(np.float32)
    elif kind == "one_front":
        x[:, 1] = 1.0 - x[:, 0]
    elif kind == "nan":
        x[::13, 2] = np.nan
    ref = C.non_dominated_sort(x)
    for prep, peel in (("", ""), ("radix", ""), ("", "dynamic"), ("radix", "dynamic")):
        monkeypatch.setenv("EVX_PREP", prep)
        monkeypatch.setenv("EVX_PEEL", peel)
        r = E.operators.non_dominated_sort(_cuda(x)).cpu().numpy()
        assert (r == ref).all(), (prep, peel, np.nonzero(r != ref)[0][:10])


@pytest.mark.parametrize("n,m", [(2600, 3), (5000, 3), (8000, 4), (2500, 2)])
def test_nds_small_population_repeatable(E, n, m):
    """The single-CTA kernels synchronise through shared memory and CTA barriers only: a missing barrier would
    show as run-to-run differences.  40 runs of one input (and the oracle once)."""
    from oracle import oracle_c as C
    rng = np.random.default_rng(n)
    x = rng.random((n, m), dtype=np.float32)
    if m == 2:
        x[:, 1] = 1.0 - x[:, 0]                                          # one front: several gather passes
    xd = _cuda(x)
    first = E.operators.non_dominated_sort(xd)
    assert (first.cpu().numpy() == C.non_dominated_sort(x)).all()
    for _ in range(40):
        assert torch.equal(E.operators.non_dominated_sort(xd), first)


@pytest.mark.parametrize("kind", ["ties", "dupes", "inf", "nan", "signed_zero", "one_front", "chain", "all_equal", "all_nan",
                                  "one_valid"])
def test_nds_edge_cases(E, kind):
    from oracle import oracle_c as C
    rng = np.random.default_rng(99)
    n = 3000
    x = rng.random((n, 3), dtype=np.float32)
    if kind == "ties":
        x = rng.integers(0, 12, (n, 3)).astype(np.float32)
    elif kind == "dupes":
        x[1000:2000] = x[:1000]
    elif kind == "inf":
        x[::5] = np.inf
        x[7, 1] = -np.inf
        x[9] = -np.inf
    elif kind == "nan":
        x[::11, 1] = np.nan
        x[5] = np.nan
    elif kind == "signed_zero":
        x = rng.integers(-1, 2, (n, 3)).astype(np.float32)
        z = x == 0
        x[z] = np.where(rng.random(int(z.sum())) < 0.5, 0.0, -0.0).astype(np.float32)
    elif kind == "one_front":
        x = (x / x.sum(1, keepdims=True)).astype(np.float32)
    elif kind == "chain":
        x = (np.arange(n, dtype=np.float32)[:, None] * np.ones((1, 3), np.float32))[rng.permutation(n)]
    elif kind == "all_equal":
        x[:] = 0.25
    elif kind == "all_nan":
        x[:, 1] = np.nan
    elif kind == "one_valid":
        x[:, 2] = np.nan
        x[1234] = 0.5
    r = E.operators.non_dominated_sort(_cuda(x)).cpu().numpy()
    ref = C.non_dominated_sort(x)
    assert (r == ref).all(), (kind, np.nonzero(r != ref)[0][:10])


def test_nds_empty(E):
    r = E.operators.non_dominated_sort(torch.empty((0, 3), device="cuda"))
    assert r.shape == (0,) and r.dtype == torch.int32


def test_nds_host_entry_point(E):
    """evx_nds_host_f32: host buffers in and out through the C ABI alone (no torch in the call)."""
    from oracle import oracle_c as C
    L = E.lib()
    x = np.random.default_rng(4).random((7000, 3), dtype=np.float32)
    rank = np.full(7000, -7, np.int32)
    rc = L.evx_nds_host_f32(x.ctypes.data_as(ctypes.c_void_p), 7000, 3, rank.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0, L.evx_last_error()
    ref = C.non_dominated_sort(x)
    assert (rank == ref).all()
    assert L.evx_nds_last_num_fronts() == int(ref.max()) + 1


def test_nds_error_reporting(E):
    L = E.lib()
    x = torch.rand(100, 3, device="cuda")
    r = torch.empty(100, dtype=torch.int32, device="cuda")
    assert L.evx_nds_f32(x.data_ptr(), 100, 3, r.data_ptr(), None, 0, None) == 2       # EVX_ERR_WORKSPACE
    assert b"workspace" in L.evx_last_error()
    assert L.evx_nds_f32(x.data_ptr(), 100, 40, r.data_ptr(), None, 0, None) == 1      # EVX_ERR_ARG


def _verify_ranks_sampled(x, rank, sample_idx):
    """Exactness check usable at any n: rank[j] == 1 + max rank over j's dominators (0 if none),
    evaluated for a sample of j against ALL rows with plain torch compares (evox_b200/_verify.py)."""
    from evox_b200._verify import rank_mismatches_sampled
    assert rank_mismatches_sampled(x, rank, sample_idx) == 0


@pytest.mark.parametrize("n,m", [(200_000, 3), (100_000, 10)])
def test_nds_full_size_properties(E, n, m):
    """BASELINE.json configs[2] (N = 2 x pop 100k, M=3) and configs[3] (N=100k, M=10)."""
    g = torch.Generator(device="cuda").manual_seed(n + m)
    x = torch.rand((n, m), device="cuda", generator=g)
    rank = E.operators.non_dominated_sort(x)
    nf = int(rank.max()) + 1
    assert (torch.bincount(rank.long(), minlength=nf) > 0).all()              # no empty front below the last
    idx = torch.randperm(n, device="cuda", generator=g)[:512]
    _verify_ranks_sampled(x, rank, idx)
    perm = torch.randperm(n, device="cuda", generator=g)                       # permutation invariance
    assert (E.operators.non_dominated_sort(x[perm]) == rank[perm]).all()
    assert (E.operators.non_dominated_sort(x) == rank).all()                   # idempotent / deterministic


def test_nds_heavy_fronts(E):
    """Three stacked copies of a non-dominated layer: every front has n/3 members and every column of
    the next layer loses n/3 dominators at once, so the peel's per-CTA share of a front exceeds what
    the register count planes hold (several accumulate passes) and most pairs are shared by CTAs."""
    n = 300_000
    g = torch.Generator(device="cuda").manual_seed(5)
    base = torch.rand((n, 3), device="cuda", generator=g)
    base = base / base.sum(1, keepdim=True)
    layer = (torch.arange(n, device="cuda") % 3).to(torch.float32)
    x = base + layer[:, None]
    rank = E.operators.non_dominated_sort(x)
    assert (rank == layer.to(torch.int32)).all()
    _verify_ranks_sampled(x, rank, torch.randperm(n, device="cuda", generator=g)[:256])


@pytest.mark.parametrize("kind", ["uniform", "stacked_simplex"])
def test_nds_1m_sampled(E, kind):
    """BASELINE.json configs[4], the headline shape: N = 1 000 000, M = 3.  `uniform` is bench.py's own
    input (numpy default_rng(0)); `stacked_simplex` is three non-dominated layers of N/3 rows (a
    converged population: few, very heavy fronts).  No dense oracle exists at this size (the reference's
    matrix would be 1 TB), so the ranks are held to the peel's defining identity on 2048 sampled rows
    against all 1M, plus "every front index is used"."""
    from evox_b200._verify import fronts_contiguous
    n = 1_000_000
    g = torch.Generator(device="cuda").manual_seed(11)
    if kind == "uniform":
        x = _cuda(np.random.default_rng(0).random((n, 3), dtype=np.float32))
    else:
        base = torch.rand((n, 3), device="cuda", generator=g)
        base = base / base.sum(1, keepdim=True)
        layer = (torch.arange(n, device="cuda") % 3).to(torch.float32)
        x = base + layer[:, None]
    rank = E.operators.non_dominated_sort(x)
    assert fronts_contiguous(rank)
    if kind == "uniform":
        assert 150 < int(rank.max()) + 1 < 400                                 # ~2.1 N^(1/3) fronts
    else:
        # float rounding can put a handful of rows a hair inside their layer, never a whole layer
        off = int((rank != (torch.arange(n, device="cuda") % 3).to(torch.int32)).sum())
        assert off < 50 and int(rank.max()) + 1 <= 6, (off, int(rank.max()))
    _verify_ranks_sampled(x, rank, torch.randperm(n, device="cuda", generator=g)[:2048])
    from evox_b200._dev import release_workspaces
    del rank
    release_workspaces()                                                       # 62.7 GB of scratch back to torch
    torch.cuda.empty_cache()


# ------------------------------------------------------------------------------------------------
# dominate_relation and the dense-matrix peel
# ------------------------------------------------------------------------------------------------
def test_relation_fixtures(E, golden):
    i = 0
    while f"rel{i}_x" in golden:
        d = E.utils.dominate_relation(_cuda(golden[f"rel{i}_x"]), _cuda(golden[f"rel{i}_y"]))
        assert d.dtype == torch.bool and (d.cpu().numpy() == golden[f"rel{i}_d"]).all(), i
        i += 1


@pytest.mark.parametrize("nx,ny,m", [(1, 1, 2), (33, 17, 3), (700, 1024, 3), (513, 2050, 4), (300, 999, 10), (64, 64, 1)])
def test_relation_random(E, nx, ny, m):
    from oracle import oracle_c as C
    rng = np.random.default_rng(nx + ny + m)
    x = rng.integers(0, 6, (nx, m)).astype(np.float32)
    y = rng.integers(0, 6, (ny, m)).astype(np.float32)
    x[0, 0] = np.nan
    ref = C.dominate_relation(x, y)
    d = E.utils.dominate_relation(_cuda(x), _cuda(y)).cpu().numpy()
    assert (d == ref).all()
    bits = E.utils.dominate_relation_bits(_cuda(x), _cuda(y)).cpu().numpy().view(np.uint32)
    unpacked = ((bits[:, :, None] >> np.arange(32, dtype=np.uint32)) & 1).reshape(nx, -1)[:, :ny].astype(bool)
    assert (unpacked == ref).all()
    if ny % 32:
        assert ((bits[:, -1] >> np.uint32(ny % 32)) == 0).all()          # tail bits are zero


@pytest.mark.parametrize("m", [2, 3, 4, 6])
def test_relation_special_values(E, m):
    """The byte-matrix kernel compares by the sign of differences for m <= 4: infinities (inf - inf), signed
    zeros and NaN in rows and in columns must come out as IEEE compares give them (utils/common.py:88-91)."""
    from oracle import oracle_c as C
    rng = np.random.default_rng(40 + m)
    vals = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 0.5, 1e-45, -1e-45], np.float32)
    x = vals[rng.integers(0, len(vals), (257, m))]
    y = vals[rng.integers(0, len(vals), (1000, m))]
    with np.errstate(invalid="ignore"):
        ref = C.dominate_relation(x, y)
    d = E.utils.dominate_relation(_cuda(x), _cuda(y)).cpu().numpy()
    assert (d == ref).all(), np.argwhere(d != ref)[:5]


def test_host_rank_from_domination_matrix(E):
    from oracle import oracle_np as O
    from evox_b200.operators.selection.non_dominate import host_rank_from_domination_matrix
    x = np.random.default_rng(8).random((900, 3), dtype=np.float32)
    dom = O.dominate_relation(x, x)
    ref = O.rank_from_domination_matrix(dom, dom.sum(0))
    r1 = host_rank_from_domination_matrix(_cuda(dom), _cuda(dom.sum(0).astype(np.int32)))
    r2 = host_rank_from_domination_matrix(E.utils.dominate_relation(_cuda(x), _cuda(x)), None)
    assert (r1.cpu().numpy() == ref).all() and (r2.cpu().numpy() == ref).all()


# ------------------------------------------------------------------------------------------------
# crowding distance
# ------------------------------------------------------------------------------------------------
def test_crowding_reference_kats(E):
    inf = np.inf
    # /root/reference/tests/test_crowding_distance.py:7-21
    x = np.array([[0.5, 4], [1, 2.5], [2, 2], [3, 1], [4, 0.8]], np.float32)
    exp = np.array([inf, 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, 2 / 3.5 + 1.2 / 3.2, inf], np.float32)
    _close(E.operators.crowding_distance(_cuda(x)).cpu().numpy(), exp, CD_RTOL)
    # :127-146
    x = np.array([[-1, -1], [0.5, 4], [1, 2.5], [2, 2], [-2, -2], [3, 1], [4, 0.8], [-3, -3], [-3, -4]], np.float32)
    mask = np.array([False, True, True, True, False, True, True, False, False])
    exp = np.array([-inf, inf, 1.5 / 3.5 + 2 / 3.2, 2 / 3.5 + 1.5 / 3.2, -inf, 2 / 3.5 + 1.2 / 3.2, inf, -inf, -inf],
                   np.float32)
    _close(E.operators.crowding_distance(_cuda(x), _cuda(mask)).cpu().numpy(), exp, CD_RTOL)


def test_crowding_prng_golden(E):
    # /root/reference/tests/test_crowding_distance.py:24-124 (input regenerated by oracle/jax_prng.py)
    from pathlib import Path
    from oracle import jax_prng
    x = jax_prng.normal(jax_prng.prng_key(314), (128, 8))
    rank = E.operators.non_dominated_sort(_cuda(x))
    pf = _cuda(x)[rank == 0]
    d = E.operators.crowding_distance(pf).cpu().numpy()
    exp = np.load(Path(__file__).parent / "golden" / "kat_crowding2.npy")
    _close(d, exp, 1e-5)                      # the reference test's own tolerance


def test_crowding_fixtures(E, golden):
    i = 0
    while f"cd{i}_c" in golden:
        c = golden[f"cd{i}_c"]
        mask = golden[f"cd{i}_mask"] if bool(golden[f"cd{i}_has_mask"]) else None
        d = E.operators.crowding_distance(_cuda(c), None if mask is None else _cuda(mask)).cpu().numpy()
        _close(d, golden[f"cd{i}_d"], CD_RTOL)
        s = E.operators.crowding_distance_sort(_cuda(c), None if mask is None else _cuda(mask)).cpu().numpy()
        assert (s == golden[f"cd{i}_sort"]).all(), i
        i += 1
    assert i >= 17


@pytest.mark.parametrize("n,m,p", [(100_000, 3, None), (200_000, 3, 0.02), (50_000, 10, 0.5), (4096, 2, 0.0)])
def test_crowding_large_vs_oracle(E, n, m, p):
    from oracle import oracle_c as C
    rng = np.random.default_rng(n + m)
    c = rng.random((n, m), dtype=np.float32)
    c[rng.integers(0, n, n // 10)] = c[rng.integers(0, n, n // 10)]           # ties
    mask = None if p is None else rng.random(n) < p
    d = E.operators.crowding_distance(_cuda(c), None if mask is None else _cuda(mask)).cpu().numpy()
    _close(d, C.crowding_distance(c, mask), CD_RTOL)


# ------------------------------------------------------------------------------------------------
# survivor selection
# ------------------------------------------------------------------------------------------------
def test_non_dominate_fixtures(E, golden):
    i = 0
    while f"sel{i}_pop" in golden:
        topk, dedup = int(golden[f"sel{i}_topk"]), bool(golden[f"sel{i}_dedup"])
        p, f = E.operators.selection.non_dominate(_cuda(golden[f"sel{i}_pop"]), _cuda(golden[f"sel{i}_fit"]), topk, dedup)
        assert (p.cpu().numpy() == golden[f"sel{i}_opop"]).all(), i
        assert (f.cpu().numpy() == golden[f"sel{i}_ofit"]).all(), i
        p2, f2 = E.operators.selection.NonDominate(topk, dedup)(_cuda(golden[f"sel{i}_pop"]), _cuda(golden[f"sel{i}_fit"]))
        assert (p2 == p).all() and (f2 == f).all()
        i += 1
    assert i >= 6


def test_non_dominate_nsga2_shape(E):
    """BASELINE.json configs[0] shape: pop 100 + 100 offspring, M=3, d=7."""
    from oracle import oracle_np as O
    rng = np.random.default_rng(21)
    pop = rng.random((200, 7), dtype=np.float32)
    fit = O.dtlz(1, pop, 3)
    p, f = E.operators.selection.non_dominate(_cuda(pop), _cuda(fit), 100, False)
    po, fo = O.non_dominate(pop, fit, 100, False)
    assert (p.cpu().numpy() == po).all() and (f.cpu().numpy() == fo).all()


# ------------------------------------------------------------------------------------------------
# problems
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("k", range(1, 8))
def test_dtlz_fixtures(E, golden, k):
    cls = getattr(E.problems.numerical, f"DTLZ{k}")
    n = 0
    for j in range(5):
        if f"dtlz{k}_{j}_X" not in golden:
            continue
        X, m = golden[f"dtlz{k}_{j}_X"], int(golden[f"dtlz{k}_{j}_m"])
        state = object()
        F, st = cls(d=X.shape[1], m=m).evaluate(state, _cuda(X))
        assert st is state
        _close(F.cpu().numpy(), golden[f"dtlz{k}_{j}_F"], EVAL_RTOL, atol=1e-6)
        n += 1
    assert n >= 4


def test_dtlz1_reference_kat(E):
    # /root/reference/tests/test_classic_problems.py:46-54
    F, _ = E.problems.numerical.DTLZ1(d=None, m=4).evaluate(None, torch.full((16, 7), 0.5, device="cuda"))
    np.testing.assert_allclose(F.sum(dim=1).cpu().numpy(), 0.5 * np.ones(16), atol=1e-6)


def test_maf1_fixtures_and_kat(E, golden):
    from oracle import jax_prng
    for j in range(3):
        X, m = golden[f"maf1_{j}_X"], int(golden[f"maf1_{j}_m"])
        F, _ = E.problems.numerical.MaF1(d=X.shape[1], m=m).evaluate(None, _cuda(X))
        _close(F.cpu().numpy(), golden[f"maf1_{j}_F"], EVAL_RTOL, atol=1e-6)
    data = jax_prng.uniform(jax_prng.prng_key(1), (3, 12))       # /root/reference/tests/test_maf.py:10-11,49-57
    F, _ = E.problems.numerical.MaF1(d=12, m=3).evaluate(None, _cuda(data))
    assert abs(float(F[1, 1]) - 1.8404) < 1e-4


_LSMOP_GOLD = {1: (12.5981, 21.2454, 0, 0), 2: (0.6876, 0.4688, 0, 0), 3: (5.3782496e04, 26.0379, 1, 0),
               4: (0.5550, 0.8865, 0, 0), 5: (2.6036, 10.7802, 0, 0), 6: (1.7095365e03, 2.5482053e04, 1, 1),
               7: (1.4722900e04, 1.3406048, 1, 0), 8: (1.9358, 0.8251, 0, 0), 9: (0.0523, 217.5438, 0, 0)}


@pytest.mark.parametrize("k", range(1, 10))
def test_lsmop_fixtures_and_kat(E, golden, k):
    from oracle import jax_prng
    cls = getattr(E.problems.numerical, f"LSMOP{k}")
    for j in range(4):
        X, m = golden[f"lsmop{k}_{j}_X"], int(golden[f"lsmop{k}_{j}_m"])
        prob = cls(d=X.shape[1], m=m)
        assert list(prob.sublen) == golden[f"lsmop{k}_{j}_sublen"].tolist()
        assert list(prob.len) == golden[f"lsmop{k}_{j}_off"].tolist()
        F, _ = prob.evaluate(None, _cuda(X))
        _close(F.cpu().numpy(), golden[f"lsmop{k}_{j}_F"], EVAL_RTOL, atol=1e-6)
    # /root/reference/tests/test_lsmop.py:10-14 and the per-problem spot values
    upper = np.concatenate([np.ones(2, np.float32), 10 * np.ones(298, np.float32)])
    data = jax_prng.uniform(jax_prng.prng_key(0), (100, 300), 0.0, upper)
    r, _ = cls(d=300, m=3).evaluate(None, _cuda(data))
    a, b, rel_a, rel_b = _LSMOP_GOLD[k]
    for got, exp, rel in ((float(r[1, 1]), a, rel_a), (float(r[1, 2]), b, rel_b)):
        assert (abs(got - exp) / exp < 1e-5) if rel else (abs(got - exp) < 1e-4), (k, got, exp)


@pytest.mark.parametrize("name,n,d,m", [("DTLZ2", 100_000, 12, 3), ("DTLZ1", 100, 7, 3), ("MaF1", 50_000, 19, 10),
                                        ("LSMOP1", 20_000, 300, 3), ("LSMOP7", 3000, 1000, 10)])
def test_problems_config_sizes_vs_oracle(E, name, n, d, m):
    """BASELINE.json configs[0], [2], [3] evaluate shapes, against the NumPy oracle."""
    from oracle import oracle_np as O
    rng = np.random.default_rng(n + d)
    X = rng.random((n, d), dtype=np.float32)
    if name.startswith("LSMOP"):
        X[:, m - 1:] *= 10
        ref = O.lsmop(int(name[5:]), X, m)
    elif name == "MaF1":
        ref = O.maf1(X, m)
    else:
        ref = O.dtlz(int(name[4:]), X, m)
    F, _ = getattr(E.problems.numerical, name)(d=d, m=m).evaluate(None, _cuda(X))
    _close(F.cpu().numpy(), ref, EVAL_RTOL, atol=1e-5)
