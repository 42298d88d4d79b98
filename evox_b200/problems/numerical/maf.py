"""MaF1 (src/evox/problems/numerical/maf.py:61-113).  MaF2-15 are not on the hot path."""
import torch

from ..._dev import as_f32_cuda, stream_ptr
from ..._lib import check, lib
from ...core import Problem


class MaF1(Problem):
    def __init__(self, d=None, m=None, ref_num=1000):
        self.m = 3 if m is None else m
        self.d = self.m + 9 if d is None else d
        self.ref_num = ref_num

    def evaluate(self, state, X):
        X = as_f32_cuda(X, "X")
        n, d = X.shape
        F = torch.empty((n, self.m), dtype=torch.float32, device="cuda")
        check(lib().evx_eval_maf1_f32(X.data_ptr(), n, d, self.m, F.data_ptr(), stream_ptr()))
        return F, state
