"""Mirror of `evox.operators.selection.non_dominate` (src/evox/operators/selection/non_dominate.py).

Same names, argument meaning and results as the reference; arrays are CUDA tensors and the work is
done by libevoxmo.so.  `method` is accepted for drop-in compatibility: the reference's three
methods ("scan", "full map-reduce", "host", :49-60) are three ways of evaluating the same ranks, and
all of them map to the one CUDA implementation.
"""
import torch

from ..._dev import as_f32_cuda, stream_ptr, workspace
from ..._lib import check, lib

_METHODS = ("auto", "scan", "full map-reduce", "host")


def non_dominated_sort(x, method="auto"):
    """Front index (int32, from 0) of every row of x [n, m].  non_dominate.py:32-115."""
    assert method in _METHODS, "method must be either 'auto', 'scan', or 'full map-reduce', 'host'"
    x = as_f32_cuda(x, "x")
    if x.dim() != 2:
        raise ValueError(f"non_dominated_sort: expected [n, m], got {tuple(x.shape)}")
    n, m = x.shape
    rank = torch.empty((n,), dtype=torch.int32, device="cuda")
    if n == 0:
        return rank
    L = lib()
    nbytes = L.evx_nds_workspace_bytes(n, m)
    if nbytes == 0:
        raise ValueError(f"non_dominated_sort: unsupported shape n={n}, m={m}")
    ws = workspace(nbytes, "nds")
    check(L.evx_nds_f32(x.data_ptr(), n, m, rank.data_ptr(), ws.data_ptr(), ws.numel(), stream_ptr()))
    return rank


def host_rank_from_domination_matrix(dominate_mat, dominate_count):
    """non_dominate.py:13-29 on the device: ranks from a dense bool matrix and its column sums."""
    d = dominate_mat if isinstance(dominate_mat, torch.Tensor) else torch.as_tensor(dominate_mat)
    d = d.to(device="cuda", dtype=torch.uint8).contiguous()
    n = d.shape[0]
    cnt = None
    if dominate_count is not None:
        cnt = (dominate_count if isinstance(dominate_count, torch.Tensor) else torch.as_tensor(dominate_count))
        cnt = cnt.to(device="cuda", dtype=torch.int32).contiguous()
    rank = torch.empty((n,), dtype=torch.int32, device="cuda")
    if n == 0:
        return rank
    L = lib()
    ws = workspace(L.evx_rank_from_matrix_workspace_bytes(n), "dense")
    check(L.evx_rank_from_matrix_u8(d.data_ptr(), None if cnt is None else cnt.data_ptr(), n, rank.data_ptr(),
                                    ws.data_ptr(), ws.numel(), stream_ptr()))
    return rank


def _mask_u8(mask, n):
    if mask is None:
        return None
    mk = mask if isinstance(mask, torch.Tensor) else torch.as_tensor(mask)
    if mk.shape != (n,):
        raise ValueError(f"mask: expected shape ({n},), got {tuple(mk.shape)}")
    return mk.to(device="cuda").ne(0).to(torch.uint8).contiguous()


def crowding_distance(costs, mask=None):
    """Crowding distance of costs [n, m]; rows with mask False get -inf.  non_dominate.py:118-158."""
    c = as_f32_cuda(costs, "costs")
    n, m = c.shape
    out = torch.empty((n,), dtype=torch.float32, device="cuda")
    if n == 0:
        return out
    mk = _mask_u8(mask, n)
    L = lib()
    ws = workspace(L.evx_crowding_workspace_bytes(n, m), "cd")
    check(L.evx_crowding_distance_f32(c.data_ptr(), None if mk is None else mk.data_ptr(), n, m, out.data_ptr(),
                                      ws.data_ptr(), ws.numel(), stream_ptr()))
    return out


def crowding_distance_sort(x, mask=None):
    """Stable ascending argsort of `crowding_distance(x, mask)`.  non_dominate.py:161-180."""
    c = as_f32_cuda(x, "x")
    n, m = c.shape
    out = torch.empty((n,), dtype=torch.int32, device="cuda")
    if n == 0:
        return out
    mk = _mask_u8(mask, n)
    L = lib()
    ws = workspace(L.evx_crowding_workspace_bytes(n, m), "cd")
    check(L.evx_crowding_distance_sort_f32(c.data_ptr(), None if mk is None else mk.data_ptr(), n, m, out.data_ptr(),
                                           ws.data_ptr(), ws.numel(), stream_ptr()))
    return out


def non_dominate(population, fitness, topk, deduplicate):
    """Survivor selection.  non_dominate.py:183-222.  Everything after the optional de-duplication
    (:210-222: non-dominated sort, cut rank, crowding permutation of the cut front, the final lexsort
    and the gather) is ONE library call, evx_non_dominate_f32, which keeps the reference's :216/:218
    key choice: the secondary key is the negated crowding *argsort permutation*.  The de-duplication
    (:188-208, `jnp.unique(axis=0, size=n)` with its index padding) stays host glue, as SURVEY 8 f1 says."""
    pop = population if isinstance(population, torch.Tensor) else torch.as_tensor(population)
    pop = pop.to("cuda")
    fit = as_f32_cuda(fitness, "fitness")
    n = pop.shape[0]
    if fit.dim() != 2 or fit.shape[0] != n:
        raise ValueError(f"non_dominate: population {tuple(pop.shape)} and fitness {tuple(fit.shape)} disagree")
    topk = int(topk)
    if not 1 <= topk <= n:
        raise ValueError(f"non_dominate: topk={topk} must be in [1, {n}]")
    fit_bak = None
    if deduplicate:
        uniq, inverse = torch.unique(pop, dim=0, return_inverse=True)
        count = uniq.shape[0]
        ar = torch.arange(n, device="cuda")
        first = torch.full((count,), n, dtype=torch.long, device="cuda").scatter_reduce(0, inverse, ar, "amin")
        uidx = torch.cat([first, first[:1].expand(n - count)])          # jnp.unique(size=n) index padding
        pop, fit = pop[uidx], fit[uidx]
        fit_bak = fit.contiguous()
        fit = torch.where((ar < count)[:, None], fit, torch.full_like(fit, float("inf"))).contiguous()
    return _select(pop, fit, fit_bak, topk)


def _select(pop, fit, fit_bak, topk, rank=None):
    """One library call: evx_non_dominate_f32 (sorts, then selects), or evx_select_by_rank_f32 when the
    caller already holds the ranks (the sharded driver: ranks come from the multi-GPU sort)."""
    n, m = fit.shape
    pop2 = pop.reshape(n, -1)
    fused_pop = pop2.dtype == torch.float32 and pop2.shape[1] > 0
    pop2 = pop2.contiguous() if fused_pop else pop2
    d = pop2.shape[1] if fused_pop else 0
    out_fit = torch.empty((topk, m), dtype=torch.float32, device="cuda")
    out_pop = torch.empty((topk, d), dtype=torch.float32, device="cuda") if fused_pop else None
    out_idx = None if fused_pop else torch.empty((topk,), dtype=torch.int32, device="cuda")
    L = lib()
    nbytes = L.evx_non_dominate_workspace_bytes(n, m)
    if nbytes == 0:
        raise ValueError(f"non_dominate: unsupported shape n={n}, m={m}")
    ws = workspace(nbytes, "sel")
    common = (n, d, m, topk, None if out_pop is None else out_pop.data_ptr(), out_fit.data_ptr(),
              None if out_idx is None else out_idx.data_ptr(), ws.data_ptr(), ws.numel(), stream_ptr())
    head = (pop2.data_ptr() if fused_pop else None, fit.data_ptr(), None if fit_bak is None else fit_bak.data_ptr())
    if rank is None:
        check(L.evx_non_dominate_f32(*head, *common))
    else:
        check(L.evx_select_by_rank_f32(*head, rank.data_ptr(), *common))
    if fused_pop:
        return out_pop.reshape((topk,) + tuple(pop.shape[1:])), out_fit
    return pop[out_idx.long()], out_fit                                  # non-float32 populations: gather by index


def select_by_rank(population, fitness, rank, topk):
    """non_dominate.py:213-222 on ranks computed elsewhere (no counterpart in the reference, whose
    selection is replicated on every device): used by evox_b200.distributed after the sharded sort."""
    pop = population.to("cuda")
    fit = as_f32_cuda(fitness, "fitness")
    rank = rank.to(device="cuda", dtype=torch.int32).contiguous()
    return _select(pop, fit, None, int(topk), rank)


class NonDominate:
    """non_dominate.py:225-232."""

    def __init__(self, topk, deduplicate=False):
        self.topk = topk
        self.deduplicate = deduplicate

    def __call__(self, population, fitness):
        return non_dominate(population, fitness, self.topk, self.deduplicate)
