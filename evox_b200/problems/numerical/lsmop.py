"""LSMOP1-9 (src/evox/problems/numerical/lsmop.py): same constructors and `evaluate(state, X)`.

Constructor quirk kept from the reference: the variable groups (`sublen`, `len`, lsmop.py:46-54) are
sized from the `d` seen by `LSMOP.__init__`, i.e. 100*m when d is None, even though LSMOP1,2,4-9
then default `self.d` to m + 4 (:116-125)."""
import ctypes

import torch

from ..._dev import as_f32_cuda, stream_ptr
from ..._lib import check, lib
from ...core import Problem


class LSMOP(Problem):
    _which = 0

    def __init__(self, d=None, m=None, ref_num=1000):
        self.nk = 5
        self.m = 3 if m is None else m
        group_d = self.m * 100 if d is None else d
        self.d = group_d
        self.ref_num = ref_num
        sub = (ctypes.c_int32 * self.m)()
        off = (ctypes.c_int32 * (self.m + 1))()
        check(lib().evx_lsmop_groups(group_d, self.m, sub, off))
        self.sublen = tuple(sub)
        self.len = tuple(off)
        self._sub, self._off = sub, off
        if d is None and self._which != 3:
            self.d = self.m + 4

    def evaluate(self, state, X):
        X = as_f32_cuda(X, "X")
        n, d = X.shape
        F = torch.empty((n, self.m), dtype=torch.float32, device="cuda")
        check(lib().evx_eval_lsmop_f32(self._which, X.data_ptr(), n, d, self.m, self._sub, self._off, F.data_ptr(),
                                       stream_ptr()))
        return F, state


def _make(k):
    return type(f"LSMOP{k}", (LSMOP,), {"_which": k})


LSMOP1, LSMOP2, LSMOP3, LSMOP4, LSMOP5, LSMOP6, LSMOP7, LSMOP8, LSMOP9 = (_make(k) for k in range(1, 10))
