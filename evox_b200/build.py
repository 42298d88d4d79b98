"""Builds evox_b200/libevoxmo.so from csrc/*.cu with nvcc for sm_100a (in-tree, no JIT cache)."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
OUT = HERE / "libevoxmo.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "-diag-suppress", "177"] + os.environ.get("EVX_NVCC_EXTRA", "").split()


def _stale(obj, deps):
    return (not obj.exists()) or any(obj.stat().st_mtime < d.stat().st_mtime for d in deps)


def build(verbose=False, force=False):
    srcs = sorted(CSRC.glob("*.cu"))
    hdrs = sorted(CSRC.glob("*.cuh")) + [HERE.parent / "include" / "evoxmo.h"]
    objdir = HERE / "build"
    objdir.mkdir(exist_ok=True)

    def one(src):
        obj = objdir / (src.stem + ".o")
        if force or _stale(obj, [src] + hdrs):
            cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)] + (["-Xptxas", "-v"] if verbose else [])
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
            if verbose:
                sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(one, srcs))
    if force or _stale(OUT, objs):
        r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", str(OUT), *map(str, objs),
                            "-lcudart"],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    build_ffi()
    return OUT


def build_ffi():
    """XLA FFI handlers (evox_b200/ffi/ffi_evoxmo.cc -> libevoxmo_ffi.so), only where JAX and its FFI
    headers exist.  Neither the build image nor the GPU box of this repo has JAX: there this is a no-op."""
    try:
        import jax
        inc = jax.ffi.include_dir()
    except Exception:
        return None
    src = HERE / "ffi" / "ffi_evoxmo.cc"
    out = HERE / "libevoxmo_ffi.so"
    if not _stale(out, [src, OUT]):
        return out
    cmd = ["g++", "-O2", "-std=c++17", "-shared", "-fPIC", f"-I{inc}", f"-I{HERE.parent / 'include'}",
           "-I/usr/local/cuda/include", str(src), "-o", str(out), f"-L{HERE}", "-levoxmo", f"-Wl,-rpath,{HERE}"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"FFI handler build failed:\n{r.stdout}\n{r.stderr}")
    return out


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
