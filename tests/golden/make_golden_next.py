"""Fixtures for the rows SURVEY.md section 8(f) marks "next", produced like reference_outputs.npz:
by EXECUTING THE REFERENCE'S OWN SOURCE (read in place from /root/reference) on the NumPy-backed jax
shim.  Run in the build container only:

    python tests/golden/make_golden_next.py

f4: SPEA2 `cal_fitness` and `_truncation` (src/evox/algorithms/mo/spea2.py:21-67).
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
import jax_shim  # noqa: E402

ref = jax_shim.install()
A = ref.asarray
spea2 = jax_shim.lift_spea2()


def main():
    out = {}
    rng = np.random.default_rng(20261107)
    cases = [rng.random((n, m), dtype=np.float32) for n, m in [(3, 2), (10, 2), (40, 3), (64, 3), (50, 5), (30, 10)]]
    cases.append(rng.integers(0, 4, (40, 3)).astype(np.float32))            # ties and duplicates
    s = rng.random((30, 3), dtype=np.float32)
    cases.append((s / s.sum(1, keepdims=True)).astype(np.float32))          # one front: raw fitness 0 everywhere
    for i, obj in enumerate(cases):
        out[f"fit{i}_obj"] = obj
        out[f"fit{i}_f"] = np.asarray(spea2.cal_fitness(A(obj))).astype(np.float32)
    for i, (n, m, k, p) in enumerate([(20, 2, 5, 1.0), (40, 3, 10, 0.7), (40, 3, 0, 0.5), (30, 2, 12, 0.5), (25, 3, 24, 1.0)]):
        obj = rng.random((n, m), dtype=np.float32)
        if i == 3:
            obj[5:10] = obj[:5]                                             # zero distances: ties broken by index
        mask = rng.random(n) < p
        k = min(k, max(0, int(mask.sum()) - 1))          # the reference's loop needs a finite distance to remove a row
        out[f"trunc{i}_obj"], out[f"trunc{i}_k"], out[f"trunc{i}_mask"] = obj, np.array(k), mask
        out[f"trunc{i}_keep"] = np.asarray(spea2.truncation(A(obj), k, A(mask))).astype(bool)
    np.savez_compressed(HERE / "next_rows_outputs.npz", **out)
    print("wrote", HERE / "next_rows_outputs.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
