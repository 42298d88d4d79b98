"""Fused survivor selection, evx_non_dominate_f32, through the C ABI: the chosen rows
and their order are compared bit-exactly with the oracle's restatement of non_dominate.py:210-222
(the six reference-run `sel*` fixtures are covered by tests/test_gpu_parity.py::test_non_dominate_fixtures
through the host mirror, which now makes this one library call)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _select(pop, fit, topk, fit_src=None, want_pop=True):
    from evox_b200._lib import check, lib
    L = lib()
    n, m = fit.shape
    d = pop.shape[1] if want_pop else 0
    nbytes = L.evx_non_dominate_workspace_bytes(n, m)
    assert nbytes > 0
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    out_pop = torch.empty((topk, d), dtype=torch.float32, device="cuda") if want_pop else None
    out_fit = torch.empty((topk, m), dtype=torch.float32, device="cuda")
    out_idx = torch.full((topk,), -1, dtype=torch.int32, device="cuda")
    check(L.evx_non_dominate_f32(pop.data_ptr() if want_pop else None, fit.data_ptr(),
                                 None if fit_src is None else fit_src.data_ptr(), n, d, m, topk,
                                 None if out_pop is None else out_pop.data_ptr(), out_fit.data_ptr(), out_idx.data_ptr(),
                                 ws.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out_pop, out_fit, out_idx


def _oracle_choice(fit, rank, topk):
    """non_dominate.py:213-218 on given ranks (NumPy)."""
    from oracle import oracle_np as O
    order = np.argsort(rank, kind="stable")
    worst = rank[order[topk - 1]]
    cd_perm = O.crowding_distance_sort(fit, rank == worst)
    return np.lexsort((-cd_perm, rank))[:topk]


def _fitness(kind, n, m, rng):
    fit = rng.random((n, m), dtype=np.float32)
    if kind == "ties":
        fit = rng.integers(0, 5, (n, m)).astype(np.float32)
    elif kind == "one_front":
        fit = (fit / fit.sum(1, keepdims=True)).astype(np.float32)
    elif kind == "chain":
        fit = (np.arange(n, dtype=np.float32)[:, None] * np.ones((1, m), np.float32))[rng.permutation(n)]
    elif kind == "flat_objective":
        fit[:, 1] = 0.5                                   # zero range: 0/0 = NaN distances, sorted last
    elif kind == "nan_inf":
        fit[::17, 0] = np.nan
        fit[5::23] = np.inf
    return fit


@pytest.mark.parametrize("n,d,m,topk,kind", [
    (200, 7, 3, 100, "random"),            # BASELINE.json configs[0]: pop 100 + 100 offspring, DTLZ1
    (200, 7, 3, 1, "random"), (200, 7, 3, 200, "random"), (301, 4, 2, 150, "random"), (1500, 12, 5, 700, "random"),
    (5000, 12, 3, 2500, "random"), (3000, 3, 3, 1500, "ties"), (4000, 5, 3, 2000, "one_front"),
    (600, 2, 2, 300, "chain"), (2000, 3, 3, 1000, "flat_objective"), (2000, 6, 3, 999, "nan_inf"),
    (1025, 1, 10, 512, "random")])
def test_select_vs_oracle(n, d, m, topk, kind):
    from oracle import oracle_np as O
    rng = np.random.default_rng(n * 13 + topk + m)
    pop = rng.random((n, d), dtype=np.float32)
    fit = _fitness(kind, n, m, rng)
    po, fo = O.non_dominate(pop, fit, topk, False)
    op, of, oi = _select(torch.from_numpy(pop).cuda(), torch.from_numpy(fit).cuda(), topk)
    idx = oi.cpu().numpy()
    assert (pop[idx] == po).all() and np.array_equal(fit[idx], fo, equal_nan=True)
    assert (op.cpu().numpy() == po).all() and np.array_equal(of.cpu().numpy(), fo, equal_nan=True)
    assert idx.min() >= 0 and len(set(idx.tolist())) == topk


@pytest.mark.parametrize("kind", ["random", "one_front"])
def test_select_config2_shape(kind):
    """BASELINE.json configs[2]: pop 100k + 100k offspring, d = 12, M = 3.  Ranks come from the library's
    own sort (held to exactness elsewhere); everything after them - cut rank, crowding permutation, the
    (rank, -permutation) order, the gather - is checked against the NumPy oracle, exactly."""
    import evox_b200 as E
    n, d, m, topk = 200_000, 12, 3, 100_000
    rng = np.random.default_rng(2)
    pop = rng.random((n, d), dtype=np.float32)
    fit = _fitness(kind, n, m, rng)
    fd = torch.from_numpy(fit).cuda()
    rank = E.operators.non_dominated_sort(fd).cpu().numpy()
    want = _oracle_choice(fit, rank, topk)
    op, of, oi = _select(torch.from_numpy(pop).cuda(), fd, topk)
    assert (oi.cpu().numpy() == want).all()
    assert (op.cpu().numpy() == pop[want]).all() and (of.cpu().numpy() == fit[want]).all()


def test_select_fitness_backup_and_index_only():
    """The de-duplicating caller ranks one fitness array and gathers another (non_dominate.py:201-208);
    populations that are not float32 ask for the chosen row numbers only."""
    from oracle import oracle_np as O
    rng = np.random.default_rng(8)
    n, m, topk = 900, 3, 400
    pop = rng.integers(0, 3, (n, 4)).astype(np.float32)             # many duplicate individuals
    fit = rng.random((n, m), dtype=np.float32)
    po, fo = O.non_dominate(pop, fit, topk, True)
    import evox_b200 as E
    p, f = E.operators.selection.non_dominate(torch.from_numpy(pop).cuda(), torch.from_numpy(fit).cuda(), topk, True)
    assert (p.cpu().numpy() == po).all() and (f.cpu().numpy() == fo).all()
    ipop = rng.integers(0, 1000, (n, 5))                             # int64 population: gathered by index
    po, fo = O.non_dominate(ipop, fit, topk, False)
    p, f = E.operators.selection.non_dominate(torch.from_numpy(ipop).cuda(), torch.from_numpy(fit).cuda(), topk, False)
    assert p.dtype == torch.int64 and (p.cpu().numpy() == po).all() and (f.cpu().numpy() == fo).all()
    _, f2, idx = _select(None, torch.from_numpy(fit).cuda(), topk, want_pop=False)
    assert (f2.cpu().numpy() == fo).all() and (ipop[idx.cpu().numpy()] == po).all()


def test_select_errors_and_async():
    from evox_b200._lib import lib
    L = lib()
    fit = torch.rand(100, 3, device="cuda")
    out = torch.empty(10, 3, device="cuda")
    ws = torch.empty(L.evx_non_dominate_workspace_bytes(100, 3), dtype=torch.uint8, device="cuda")
    args = (None, fit.data_ptr(), None, 100, 0, 3)
    assert L.evx_non_dominate_f32(*args, 0, None, out.data_ptr(), None, ws.data_ptr(), ws.numel(), None) == 1     # topk < 1
    assert L.evx_non_dominate_f32(*args, 101, None, out.data_ptr(), None, ws.data_ptr(), ws.numel(), None) == 1   # topk > n
    assert b"topk" in L.evx_last_error()
    assert L.evx_non_dominate_f32(*args, 10, None, out.data_ptr(), None, ws.data_ptr(), 16, None) == 2            # workspace
    assert L.evx_non_dominate_f32(*args, 10, None, out.data_ptr(), None, ws.data_ptr(), ws.numel(), None) == 0
    torch.cuda.synchronize()
    assert L.evx_nds_last_num_fronts() >= 1
