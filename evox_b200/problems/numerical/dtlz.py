"""DTLZ1-7 (src/evox/problems/numerical/dtlz.py): same constructors and `evaluate(state, X)`;
the arithmetic is one fused CUDA kernel per call (csrc/problems.cu).  `pf()` (reference fronts for
metrics) is outside the hot path and not mirrored."""
import torch

from ..._dev import as_f32_cuda, stream_ptr
from ..._lib import check, lib
from ...core import Problem


class DTLZTestSuit(Problem):
    _which = 0
    _default_extra = 9          # d defaults to m + 9 (dtlz.py:77-80), DTLZ1: m + 4, DTLZ7: m + 19

    def __init__(self, d=None, m=None, ref_num=1000):
        self.m = 3 if m is None else m
        self.d = self.m + self._default_extra if d is None else d
        self.ref_num = ref_num

    def evaluate(self, state, X):
        X = as_f32_cuda(X, "X")
        if X.dim() != 2:
            raise ValueError(f"evaluate: expected [n, d], got {tuple(X.shape)}")
        n, d = X.shape
        F = torch.empty((n, self.m), dtype=torch.float32, device="cuda")
        check(lib().evx_eval_dtlz_f32(self._which, X.data_ptr(), n, d, self.m, F.data_ptr(), stream_ptr()))
        return F, state


class DTLZ1(DTLZTestSuit):
    _which, _default_extra = 1, 4

    def __init__(self, d=None, m=None, ref_num=100):
        super().__init__(d, m, ref_num)


class DTLZ2(DTLZTestSuit):
    _which = 2


class DTLZ3(DTLZTestSuit):
    _which = 3


class DTLZ4(DTLZTestSuit):
    _which = 4


class DTLZ5(DTLZTestSuit):
    _which = 5


class DTLZ6(DTLZTestSuit):
    _which = 6


class DTLZ7(DTLZTestSuit):
    _which, _default_extra = 7, 19
