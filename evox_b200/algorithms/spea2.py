"""SPEA2's fitness assignment and archive truncation on the CUDA hot path (mirror of the two jitted
helpers of src/evox/algorithms/mo/spea2.py:21-67; the algorithm class around them stays a caller)."""
import torch

from .._dev import as_f32_cuda, stream_ptr, workspace
from .._lib import check, lib


def cal_fitness(obj):
    """SPEA2 fitness = 1 / (distance to the second nearest row + 2) + sum of the strengths of the row's
    dominators.  spea2.py:21-37.  obj [n, m] -> float32 [n]."""
    o = as_f32_cuda(obj, "obj")
    n, m = o.shape
    out = torch.empty((n,), dtype=torch.float32, device="cuda")
    if n == 0:
        return out
    L = lib()
    nbytes = L.evx_spea2_workspace_bytes(n, m)
    if nbytes == 0:
        raise ValueError(f"cal_fitness: unsupported shape n={n}, m={m}")
    ws = workspace(nbytes, "spea2")
    check(L.evx_spea2_fitness_f32(o.data_ptr(), n, m, out.data_ptr(), ws.data_ptr(), ws.numel(), stream_ptr()))
    return out


def _truncation(obj, k, mask):
    """k times remove the valid row whose nearest valid neighbour is closest (ties: lowest index);
    returns the bool mask of the rows kept.  spea2.py:40-67.  As in the reference, k must stay below the
    number of valid rows (its while_loop does not terminate otherwise)."""
    o = as_f32_cuda(obj, "obj")
    n, m = o.shape
    mk = (mask if isinstance(mask, torch.Tensor) else torch.as_tensor(mask)).to("cuda").ne(0).to(torch.uint8).contiguous()
    if mk.shape != (n,):
        raise ValueError(f"_truncation: mask shape {tuple(mk.shape)} != ({n},)")
    keep = torch.empty((n,), dtype=torch.uint8, device="cuda")
    if n == 0:
        return keep.bool()
    k = int(k)
    L = lib()
    ws = workspace(L.evx_spea2_workspace_bytes(n, m), "spea2")
    check(L.evx_spea2_truncation_f32(o.data_ptr(), n, m, k, mk.data_ptr(), keep.data_ptr(), ws.data_ptr(), ws.numel(),
                                     stream_ptr()))
    return keep.bool()
