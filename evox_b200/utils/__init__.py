"""Mirror of `evox.utils` for the hot path: `dominate_relation` (src/evox/utils/common.py:94-97)."""
import torch

from .._dev import as_f32_cuda, stream_ptr
from .._lib import check, lib


def dominate_relation(x, y):
    """A[i, j] is True iff x[i] dominates y[j] (all <=, any <).  Returns a bool [nx, ny] CUDA tensor."""
    x = as_f32_cuda(x, "x")
    y = as_f32_cuda(y, "y")
    if x.dim() != 2 or y.dim() != 2 or x.shape[1] != y.shape[1]:
        raise ValueError(f"dominate_relation: expected [nx, m] and [ny, m], got {tuple(x.shape)} and {tuple(y.shape)}")
    nx, m = x.shape
    ny = y.shape[0]
    out = torch.empty((nx, ny), dtype=torch.bool, device="cuda")
    check(lib().evx_dominate_relation_u8(x.data_ptr(), nx, y.data_ptr(), ny, m, out.data_ptr(), stream_ptr()))
    return out


def dominate_relation_bits(x, y):
    """Bit-packed variant: int32 [nx, ceil(ny/32)], bit j%32 of word j//32 is pair (i, j)."""
    x = as_f32_cuda(x, "x")
    y = as_f32_cuda(y, "y")
    nx, m = x.shape
    ny = y.shape[0]
    out = torch.empty((nx, (ny + 31) // 32), dtype=torch.int32, device="cuda")
    check(lib().evx_dominate_relation_bits(x.data_ptr(), nx, y.data_ptr(), ny, m, out.data_ptr(), stream_ptr()))
    return out
