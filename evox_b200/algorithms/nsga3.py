"""NSGA-III's environmental selection on the CUDA hot path (mirror of NSGA3.tell,
src/evox/algorithms/mo/nsga3.py:94-210).  The sort, the association with the reference directions and
the niching loop run in libevoxmo.so; the normalisation between them (:103-139: ideal point, extreme
points, an M x M solve) is host glue in torch, as it is jnp glue in the reference."""
import torch

from .._dev import as_f32_cuda, stream_ptr, workspace
from .._lib import check, lib
from ..operators.selection.non_dominate import non_dominated_sort


def associate(normalized, ref):
    """nsga3.py:140-147 -> (group_id int32 [n], group_dist float32 [n])."""
    x = as_f32_cuda(normalized, "normalized")
    r = as_f32_cuda(ref, "ref")
    n, m = x.shape
    gid = torch.empty((n,), dtype=torch.int32, device="cuda")
    gd = torch.empty((n,), dtype=torch.float32, device="cuda")
    check(lib().evx_nsga3_associate_f32(x.data_ptr(), r.data_ptr(), n, r.shape[0], m, gid.data_ptr(), gd.data_ptr(), stream_ptr()))
    return gid, gd


def niche(rank, last_rank, group_id, group_dist, nref, pop_size, rand_words):
    """nsga3.py:148-205 -> ascending indices of the pop_size survivors.  rand_words: uint32/int32
    [pop_size + 1, 2], the two words `jax.random.randint(keys[i], ...)` consumes for loop index i."""
    rank = rank.to(device="cuda", dtype=torch.int32).contiguous()
    gid = group_id.to(device="cuda", dtype=torch.int32).contiguous()
    gd = as_f32_cuda(group_dist, "group_dist")
    rw = rand_words.to(device="cuda").contiguous().view(torch.int32).reshape(-1)
    n = rank.shape[0]
    if rw.numel() < 2 * (pop_size + 1):
        raise ValueError("niche: rand_words needs 2 * (pop_size + 1) words")
    out = torch.empty((pop_size,), dtype=torch.int32, device="cuda")
    L = lib()
    nbytes = L.evx_nsga3_niche_workspace_bytes(n, nref)
    if nbytes == 0:
        raise ValueError(f"niche: unsupported shape n={n}, nref={nref}")
    ws = workspace(nbytes, "nsga3")
    check(L.evx_nsga3_niche(rank.data_ptr(), int(last_rank), gid.data_ptr(), gd.data_ptr(), n, nref, pop_size, rw.data_ptr(),
                            out.data_ptr(), ws.data_ptr(), ws.numel(), stream_ptr()))
    return out


def normalize(merged_fitness, rank, last_rank):
    """nsga3.py:103-139 in torch (float32)."""
    fit = as_f32_cuda(merged_fitness, "fitness")
    m = fit.shape[1]
    nan = torch.full_like(fit, float("nan"))
    ranked = torch.where((rank <= last_rank)[:, None], fit, nan)
    big = torch.full_like(fit, float("inf"))
    ideal = torch.where(ranked.isnan(), big, ranked).amin(dim=0)                               # nanmin
    ranked = ranked - ideal
    weight = torch.eye(m, device="cuda") + 1e-6
    ext = []
    for i in range(m):
        v = ranked / weight[i]
        mx = torch.where(v.isnan(), -big, v).amax(dim=1)                                       # nanmax per row
        mx = torch.where(ranked.isnan().all(dim=1), big[:, 0], mx)                             # all-NaN rows never win
        ext.append(torch.argmin(mx))
    extreme = ranked[torch.stack(ext)]
    if int(torch.linalg.matrix_rank(extreme)) == m:
        plane = torch.linalg.solve(extreme, torch.ones(m, device="cuda"))
        nadir = 1.0 / plane
    else:
        nadir = torch.where(ranked.isnan(), -big, ranked).amax(dim=0)
    return ranked / nadir


def environmental_selection(merged_pop, merged_fitness, ref, pop_size, rand_words):
    """NSGA3.tell on a merged population: returns (population, fitness) of the survivors."""
    fit = as_f32_cuda(merged_fitness, "fitness")
    rank = non_dominated_sort(fit)
    order = torch.sort(rank, stable=True).indices
    last_rank = int(rank[order[pop_size]])                                                     # :101 (index pop_size)
    normalized = normalize(fit, rank, last_rank)
    gid, gd = associate(normalized, ref)
    sel = niche(rank, last_rank, gid, gd, ref.shape[0], pop_size, rand_words).long()
    pop = merged_pop.to("cuda")
    return pop[sel], fit[sel]
