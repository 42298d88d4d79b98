"""evx_nds_noalloc_f32: the non-dominated sort for callers that cannot allocate device memory.

Some callers of this library only ever hold their input and output buffers - a custom-call handler
inside another framework, a callback that runs while that framework's allocator is locked - and so
cannot pass the workspace evx_nds_f32 asks for.  This entry point keeps its scratch in a fixed array
inside the library instead.  The test calls it through the C ABI and checks its ranks against
evx_nds_f32 and against the CPU oracle (exact, as for every rank comparison in this suite), twice in
a row on the same device:
the second call runs on a smaller, different population, so anything left over from the first call in
the shared scratch would show up as a wrong rank.  The file name sorts last on purpose.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _reference_ranks(L, x):
    n, m = x.shape
    nbytes = L.evx_nds_workspace_bytes(n, m)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    rank = torch.empty(n, dtype=torch.int32, device="cuda")
    rc = L.evx_nds_f32(x.data_ptr(), n, m, rank.data_ptr(), ws.data_ptr(), nbytes, torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.evx_last_error()
    return rank


def _noalloc_ranks(L, x):
    n, m = x.shape
    rank = torch.full((n,), -5, dtype=torch.int32, device="cuda")
    rc = L.evx_nds_noalloc_f32(x.data_ptr(), n, m, rank.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == 0, L.evx_last_error()
    torch.cuda.synchronize()
    return rank


def test_noalloc_matches_workspace_entry_point_twice():
    from evox_b200._lib import lib
    from oracle import oracle_c as C
    L = lib()
    rng = np.random.default_rng(2024)
    first = rng.random((4096, 3), dtype=np.float32)           # the documented maximum
    first[100:200] = first[:100]                               # duplicates
    first[7, 1] = np.nan                                       # a NaN row ranks 0
    second = rng.integers(0, 9, (1500, 4)).astype(np.float32)  # smaller, many ties, other m
    for host in (first, second):
        x = torch.from_numpy(host).cuda()
        got = _noalloc_ranks(L, x)
        want = _reference_ranks(L, x)
        assert torch.equal(got, want), (host.shape, (got != want).nonzero()[:5])
        # ... and both agree with the dense-matrix oracle, so the two kernels cannot be wrong together
        assert (got.cpu().numpy() == C.non_dominated_sort(host)).all(), host.shape


def test_noalloc_rejects_oversize_and_allocates_nothing():
    from evox_b200._lib import lib
    L = lib()
    x = torch.rand(4097, 3, device="cuda")
    rank = torch.empty(4097, dtype=torch.int32, device="cuda")
    assert L.evx_nds_noalloc_f32(x.data_ptr(), 4097, 3, rank.data_ptr(), None) == 1      # EVX_ERR_ARG
    assert b"4096" in L.evx_last_error()
    # no device memory appears or disappears across calls (driver-level view, not torch's cache)
    x = torch.rand(3000, 3, device="cuda")
    rank = torch.empty(3000, dtype=torch.int32, device="cuda")
    _noalloc_ranks(L, x)
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(5):
        assert L.evx_nds_noalloc_f32(x.data_ptr(), 3000, 3, rank.data_ptr(), None) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free0 == free1
