"""ctypes loader for oracle/oracle_c.c (test infrastructure; see that file's header)."""
import ctypes
import hashlib
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB = None


def build():
    """Portable build (no -march): this .so travels from the build container to the GPU box."""
    subprocess.run(["make", "-s", "-C", str(_HERE)], check=True)


def _native_so():
    """A -march=native build keyed by the host CPU model, made on first use where gcc exists, so
    the timed CPU baseline is not handicapped by the portable build's SSE2 code."""
    try:
        model = next(l for l in open("/proc/cpuinfo") if l.startswith("model name")).split(":", 1)[1].strip()
    except Exception:
        return None
    tag = hashlib.sha1(model.encode()).hexdigest()[:10]
    so = _HERE / f"liboracle_c.native-{tag}.so"
    if not so.exists():
        try:
            subprocess.run(["gcc", "-O3", "-march=native", "-pthread", "-fno-fast-math", "-shared", "-fPIC",
                            "-o", str(so), str(_HERE / "oracle_c.c"), "-lm"], check=True, capture_output=True)
        except Exception:
            return None
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = _native_so()
        if so is None:
            so = _HERE / "liboracle_c.so"
            if not so.exists():
                build()
        L = ctypes.CDLL(str(so))
        L.orc_nds.restype = ctypes.c_int
        L.orc_nds.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]
        L.orc_dominate_relation.restype = None
        L.orc_dominate_relation.argtypes = [ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int64,
                                            ctypes.c_int, ctypes.c_void_p]
        L.orc_crowding.restype = None
        L.orc_crowding.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_void_p]
        L.orc_num_threads.restype = ctypes.c_int
        _LIB = L
    return _LIB


def num_threads():
    return lib().orc_num_threads()


def non_dominated_sort(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    n, m = x.shape
    rank = np.zeros(n, np.int32)
    if lib().orc_nds(x.ctypes.data, n, m, rank.ctypes.data) < 0:
        raise MemoryError("oracle N x N matrix")
    return rank


def dominate_relation(x, y):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.ascontiguousarray(y, dtype=np.float32)
    out = np.zeros((x.shape[0], y.shape[0]), np.uint8)
    lib().orc_dominate_relation(x.ctypes.data, x.shape[0], y.ctypes.data, y.shape[0], x.shape[1], out.ctypes.data)
    return out.astype(bool)


def crowding_distance(costs, mask=None):
    c = np.ascontiguousarray(costs, dtype=np.float32)
    n, m = c.shape
    out = np.zeros(n, np.float32)
    mk = None if mask is None else np.ascontiguousarray(mask, dtype=np.uint8)
    lib().orc_crowding(c.ctypes.data, None if mk is None else mk.ctypes.data, n, m, out.ctypes.data)
    return out
