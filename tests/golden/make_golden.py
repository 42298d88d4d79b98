"""Generate tests/golden/*.npz by EXECUTING THE REFERENCE'S OWN SOURCE FILES
(read in place from /root/reference, on the NumPy-backed jax shim in jax_shim.py)
on seeded inputs.  Run in the build container only:

    python tests/golden/make_golden.py

The fixtures are small (a few hundred rows) and committed; nothing at test time
reads /root/reference.
"""
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE))
sys.path.insert(0, str(HERE.parent.parent / "oracle"))
import jax_shim  # noqa: E402

ref = jax_shim.install()
A = ref.asarray


def nds_inputs():
    rng = np.random.default_rng(20261106)
    cases = []
    for n, m in [(1, 3), (2, 2), (11, 2), (37, 3), (64, 1), (100, 3), (200, 3), (200, 2),
                 (150, 5), (120, 10), (300, 3)]:
        cases.append(rng.random((n, m), dtype=np.float32))
    # heavy ties / duplicates
    cases.append(rng.integers(0, 4, (150, 3)).astype(np.float32))
    cases.append(rng.integers(0, 3, (90, 2)).astype(np.float32))
    # one front (simplex) and a chain
    s = rng.random((80, 3), dtype=np.float32)
    cases.append((s / s.sum(1, keepdims=True)).astype(np.float32))
    cases.append((np.arange(40, dtype=np.float32)[:, None] * np.ones((1, 3), np.float32)))
    # +inf rows (dedupe path produces them), -inf, NaN, signed zeros
    x = rng.random((120, 3), dtype=np.float32)
    x[::7] = np.inf
    x[3, 1] = -np.inf
    cases.append(x)
    x = rng.random((100, 3), dtype=np.float32)
    x[5, 0] = np.nan
    x[17] = np.nan
    x[40, 2] = np.nan
    cases.append(x)
    x = rng.integers(-1, 2, (60, 2)).astype(np.float32)
    x[x == 0] = np.where(rng.random(np.count_nonzero(x == 0)) < 0.5, 0.0, -0.0)
    cases.append(x.astype(np.float32))
    # integer-valued fitness as the monitors pass it (tests/test_monitors.py:34)
    cases.append(np.array([[1, 2], [3, 1], [5, 6], [0.5, 1.5], [7, 8], [0, 10]], np.float32))
    return cases


def main():
    out = {}
    # ---- non_dominated_sort + dominate_relation -------------------------------------
    for i, x in enumerate(nds_inputs()):
        r_auto = np.asarray(ref.nd.non_dominated_sort(A(x)))
        r_host = np.asarray(ref.nd.non_dominated_sort(A(x), "host"))
        r_full = np.asarray(ref.nd.non_dominated_sort(A(x), "full map-reduce"))
        assert (r_auto == r_host).all() and (r_auto == r_full).all()
        out[f"nds{i}_x"] = x
        out[f"nds{i}_rank"] = r_auto.astype(np.int32)
    rng = np.random.default_rng(7)
    for i, (nx, ny, m) in enumerate([(13, 29, 3), (50, 7, 2), (33, 65, 10), (40, 40, 4)]):
        x = rng.integers(0, 5, (nx, m)).astype(np.float32) if i == 3 else rng.random((nx, m), dtype=np.float32)
        y = rng.integers(0, 5, (ny, m)).astype(np.float32) if i == 3 else rng.random((ny, m), dtype=np.float32)
        out[f"rel{i}_x"], out[f"rel{i}_y"] = x, y
        out[f"rel{i}_d"] = np.asarray(ref.utils.dominate_relation(A(x), A(y))).astype(bool)

    # ---- crowding distance (+ sort) ---------------------------------------------------
    rng = np.random.default_rng(11)
    cd_cases = []
    for n, m in [(1, 2), (2, 3), (3, 2), (50, 2), (128, 3), (200, 8), (97, 10)]:
        cd_cases.append((rng.random((n, m), dtype=np.float32), None))
    for n, m, p in [(60, 3, 0.5), (200, 3, 0.1), (40, 2, 0.0), (40, 2, 1.0), (30, 3, 0.04), (30, 3, 0.07)]:
        cd_cases.append((rng.random((n, m), dtype=np.float32), rng.random(n) < p))
    t = rng.integers(0, 5, (80, 3)).astype(np.float32)       # ties incl. zero range in col 2
    t[:, 2] = 1.0
    cd_cases.append((t, None))
    cd_cases.append((t, rng.random(80) < 0.6))
    m1 = np.zeros(25, bool); m1[9] = True                    # exactly one valid
    cd_cases.append((rng.random((25, 3), dtype=np.float32), m1))
    m2 = np.zeros(25, bool); m2[[4, 20]] = True              # exactly two valid
    cd_cases.append((rng.random((25, 3), dtype=np.float32), m2))
    for i, (c, mk) in enumerate(cd_cases):
        d = ref.nd.crowding_distance(A(c)) if mk is None else ref.nd.crowding_distance(A(c), A(mk))
        s = ref.nd.crowding_distance_sort(A(c)) if mk is None else ref.nd.crowding_distance_sort(A(c), A(mk))
        out[f"cd{i}_c"] = c
        out[f"cd{i}_mask"] = np.zeros(0, bool) if mk is None else mk
        out[f"cd{i}_has_mask"] = np.array(mk is not None)
        out[f"cd{i}_d"] = np.asarray(d).astype(np.float32)
        out[f"cd{i}_sort"] = np.asarray(s).astype(np.int32)

    # ---- non_dominate (survivor selection) ------------------------------------------------
    rng = np.random.default_rng(13)
    for i, (n, d, m, topk, dedup) in enumerate([(40, 5, 3, 20, False), (200, 7, 3, 100, False),
                                                 (100, 4, 2, 37, False), (60, 3, 3, 30, True),
                                                 (50, 6, 5, 50, False), (80, 4, 3, 1, False)]):
        pop = rng.random((n, d), dtype=np.float32)
        fit = rng.random((n, m), dtype=np.float32)
        if dedup:
            pop[10:25] = pop[:15]
            fit[10:25] = fit[:15]
        p, f = ref.nd.non_dominate(A(pop), A(fit), topk, dedup)
        out[f"sel{i}_pop"], out[f"sel{i}_fit"] = pop, fit
        out[f"sel{i}_topk"], out[f"sel{i}_dedup"] = np.array(topk), np.array(dedup)
        out[f"sel{i}_opop"], out[f"sel{i}_ofit"] = np.asarray(p), np.asarray(f)

    # ---- problems -----------------------------------------------------------------------------
    rng = np.random.default_rng(17)
    for k in range(1, 8):
        for j, (n, d, m) in enumerate([(33, 7, 3), (20, 12, 3), (16, 14, 5), (9, 19, 10), (5, 4, 2)]):
            cls = getattr(ref.dtlz, f"DTLZ{k}")
            if k in (5, 6) and m < 3:
                continue
            X = rng.random((n, d), dtype=np.float32)
            F, _ = cls(d=d, m=m).evaluate(None, A(X))
            out[f"dtlz{k}_{j}_X"], out[f"dtlz{k}_{j}_m"], out[f"dtlz{k}_{j}_F"] = X, np.array(m), np.asarray(F)
    for j, (n, d, m) in enumerate([(12, 19, 10), (30, 12, 3), (7, 6, 4)]):
        X = rng.random((n, d), dtype=np.float32)
        F, _ = ref.maf.MaF1(d=d, m=m).evaluate(None, A(X))
        out[f"maf1_{j}_X"], out[f"maf1_{j}_m"], out[f"maf1_{j}_F"] = X, np.array(m), np.asarray(F)
    for k in range(1, 10):
        for j, (n, d, m) in enumerate([(10, 300, 3), (6, 120, 2), (5, 500, 5), (4, 77, 3)]):
            X = rng.random((n, d), dtype=np.float32)
            X[:, m - 1:] *= 10
            prob = getattr(ref.lsmop, f"LSMOP{k}")(d=d, m=m)
            F, _ = prob.evaluate(None, A(X))
            out[f"lsmop{k}_{j}_X"], out[f"lsmop{k}_{j}_m"], out[f"lsmop{k}_{j}_F"] = X, np.array(m), np.asarray(F)
            out[f"lsmop{k}_{j}_sublen"] = np.array(prob.sublen, np.int32)
            out[f"lsmop{k}_{j}_off"] = np.array(prob.len, np.int32)

    np.savez_compressed(HERE / "reference_outputs.npz", **out)
    print("wrote", HERE / "reference_outputs.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
