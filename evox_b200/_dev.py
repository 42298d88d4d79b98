"""Device-array plumbing for the host layer: torch supplies device memory and streams only."""
import torch

from ._lib import EvoxMoError


def require_cuda():
    if not torch.cuda.is_available():
        raise EvoxMoError("evox_b200 needs a CUDA device: the kernels are sm_100a only and there is no CPU path")


def as_f32_cuda(x, name="x"):
    """Accept torch tensors (any device) or array-likes; return a contiguous float32 CUDA tensor.
    float64 / float16 / bfloat16 inputs are rounded to float32, exactly what `jnp.asarray` does to a
    NumPy float64 array in the reference (x64 is never enabled there), so `non_dominated_sort(
    np.random.rand(n, 3))` works as it does upstream.  Integers are cast when the cast keeps every
    comparison exact (|int| < 2**24; the reference compares integers as integers,
    tests/test_monitors.py:34) and rejected otherwise."""
    require_cuda()
    t = x if isinstance(x, torch.Tensor) else torch.as_tensor(x)
    if not t.is_floating_point() and t.dtype != torch.bool:
        if t.numel() and int(t.abs().max()) >= (1 << 24):
            raise EvoxMoError(f"{name}: integer magnitudes >= 2**24 are not exactly representable in float32")
    return t.to(device="cuda", dtype=torch.float32).contiguous()


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


_WS = {}


def workspace(nbytes, tag="ws"):
    """Grow-only per-device scratch tensor (torch caching allocator; no cudaMalloc per call)."""
    dev = torch.cuda.current_device()
    key = (dev, tag)
    t = _WS.get(key)
    if t is None or t.numel() < nbytes:
        _WS[key] = None
        t = torch.empty(int(nbytes) + 256, dtype=torch.uint8, device="cuda")
        _WS[key] = t
    return t


def release_workspaces():
    _WS.clear()
