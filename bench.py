#!/usr/bin/env python
"""bench.py - ND-sort pairs/s on synthetic random populations (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--n 1000000] [--m 3] [--impl reference]

One "step" is one full non_dominated_sort of an [n, m] float32 population (sort + unique, dominance
tiles into the packed bit matrix, front peeling, ranks).  At N GPUs the bit matrix is sharded by
column panels (evox_b200/distributed.py) and n stays fixed: strong scaling.  One JSON line on stdout.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

# ncu figures for the kernels at HEAD (DRAM bytes per launch at N = 1M, M = 3; issue-slot utilisation and
# warp instructions per tested pair) live in this file, written from the ncu captures under profiles/
NCU_SUMMARY = ROOT / "profiles" / "r2_ncu_summary.json"
GATHER_PEAK_256B = 6760.0      # GB/s, random 256-byte rows through the peel's own ring (profiles/r1_gather_probe.txt)
METRIC = "nd_sort_pairs_per_s"
UNIT = "pairs/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", "--rows", dest="n", type=int, default=1_000_000)
    ap.add_argument("--m", type=int, default=3)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--dist", default="uniform", choices=["uniform", "one_front", "chain"],
                    help="input distribution (SURVEY 8d): i.i.d. uniform, one front (points on the simplex, a converged "
                         "population), chain (n fronts of one row; n <= 20000)")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the side measurements (other seeds, one-front inputs, NSGA-II, operators)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-sample", type=int, default=0, help="rows of the CPU-baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="fused", choices=["fused", "nccl"],
                    help="multi-GPU front exchange: in-kernel peer stores (default) or one NCCL all-gather per front")
    ap.add_argument("--no-nsga2", action="store_true", help="skip the NSGA-II generations/s side measurement")
    ap.add_argument("--verify-rows", type=int, default=2048,
                    help="rows of the sampled exactness check run on the ranks of the last timed step (outside the timing)")
    return ap.parse_args()


def workload_name(n, m, dist="uniform"):
    kind = {"uniform": "uniform-random", "one_front": "one-front (simplex)", "chain": "chain"}[dist]
    return f"non_dominated_sort N={n} M={m} {kind} float32 (BASELINE.json configs[4] shape)"


def make_input(n, m, seed, dist="uniform"):
    """Synthetic objectives.  uniform: rng.random((n, m)); one_front: the same rows scaled onto
    the simplex sum(f) = 1 (a converged population: one front, up to rounding); chain: f_i = (i, .., i),
    shuffled (n fronts of one row each)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    x = rng.random((n, m), dtype=np.float32)
    if dist == "one_front":
        x = (x / x.sum(1, keepdims=True)).astype(np.float32)
    elif dist == "chain":
        assert n <= 20000, "chain has n fronts: small n only"
        x = (np.arange(n, dtype=np.float32)[:, None] * np.ones((1, m), np.float32))[rng.permutation(n)]
    return x


# ------------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm (oracle port; the reference itself is Python/JAX and JAX is not installable here)
# ------------------------------------------------------------------------------------------------
def cpu_oracle_rate(n_rows, m, seed, reps=1):
    import numpy as np
    from oracle import oracle_c as C
    x = np.random.default_rng(seed).random((n_rows, m), dtype=np.float32)
    t0 = time.perf_counter()
    for _ in range(reps):
        C.non_dominated_sort(x)
    dt = (time.perf_counter() - t0) / reps
    return float(n_rows) ** 2 / dt, dt, C.num_threads()


def auto_cpu_sample(m, seed, target_s=12.0, user=0):
    if user:
        return user
    rate, dt, _ = cpu_oracle_rate(8000, m, seed)
    n = int((rate * target_s) ** 0.5)
    return max(8000, min(n, 60000))          # <= 3.6 GB byte matrix, as the reference materialises it


def run_reference(args):
    """--impl reference: the reference's CPU path for this metric.  EvoX is pure Python on JAX and
    JAX cannot be installed in this image (no index), so this arm times the C restatement of its
    algorithm (oracle/oracle_c.c: dense N x N byte matrix + per-front peel), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ns = auto_cpu_sample(args.m, args.seed, target_s=8.0, user=args.cpu_sample)
    for _ in range(min(args.warmup, 1)):
        cpu_oracle_rate(min(ns, 8000), args.m, args.seed)
    t0 = time.perf_counter()
    cores = 1
    for s in range(args.steps):
        _, _, cores = cpu_oracle_rate(ns, args.m, args.seed + s)
    dt = time.perf_counter() - t0
    value = args.steps * float(ns) ** 2 / dt
    sample = f"non_dominated_sort of {ns} rows per step (N x N byte matrix as the reference builds it); " \
             f"pairs/s is size-independent for the O(N^2) path"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.n, args.m), "sample_rows": ns},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def verify_last_step(torch, dist, x, ranks, world, rank, rows):
    """Outside the timed region: the ranks the LAST timed step produced are held to the peel's defining
    identity (evox_b200/_verify.py) on `rows` sampled rows against all n, and to "every front index is
    used".  With several GPUs every rank checks its own share of the sample on its own copy of the result
    and the verdicts are all-reduced, so a rank that ended with different ranks fails the run."""
    from evox_b200._verify import fronts_contiguous, rank_mismatches_sampled
    n = x.shape[0]
    g = torch.Generator(device="cuda").manual_seed(12345)
    idx = torch.randperm(n, device="cuda", generator=g)[:min(rows, n)]
    bad = rank_mismatches_sampled(x, ranks, idx[rank::world]) + (0 if fronts_contiguous(ranks) else 1)
    t = torch.tensor([bad, int(ranks.long().sum() % 1000003)], device="cuda", dtype=torch.int64)
    same = True
    if world > 1:
        lo, hi = t[1:].clone(), t[1:].clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool((lo == hi).all())                    # every rank holds the same rank vector
        dist.all_reduce(t[:1], op=dist.ReduceOp.SUM)
    return {"rows": int(idx.numel()), "mismatches": int(t[0]), "ranks_identical_on_all_gpus": same,
            "ok": int(t[0]) == 0 and same, "how": "rank[j] == 1 + max rank of j's dominators, sampled j vs all rows, plain torch"}


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import evox_b200
    from evox_b200._lib import check, lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    L = lib()
    n, m, K, W = args.n, args.m, args.steps, max(args.warmup, 3)
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    x_host = torch.from_numpy(make_input(n, m, args.seed, args.dist)).pin_memory()
    x = x_host.cuda(non_blocking=True)
    stream = torch.cuda.current_stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        for _ in range(steps):
            fn()
        e1.record(stream)
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms)

    stages = {}
    if world == 1:
        nbytes = L.evx_nds_workspace_bytes(n, m)
        ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        rank_out = torch.empty(n, dtype=torch.int32, device="cuda")
        acc = {"prep": 0.0, "tiles": 0.0, "peel": 0.0, "calls": 0}

        def step():
            check(L.evx_nds_f32(x.data_ptr(), n, m, rank_out.data_ptr(), ws.data_ptr(), ws.numel(), stream.cuda_stream))
            if acc["on"]:
                a, b, c = ctypes.c_float(), ctypes.c_float(), ctypes.c_float()
                if L.evx_profile_last(ctypes.byref(a), ctypes.byref(b), ctypes.byref(c)) == 0:
                    acc["prep"] += a.value; acc["tiles"] += b.value; acc["peel"] += c.value; acc["calls"] += 1
        acc["on"] = False
        for _ in range(W):
            step()
        sampler = ClockSampler(local)
        sampler.start()
        L.evx_launch_count_reset()
        L.evx_profile_enable(1)
        acc["on"] = True
        ms = timed(step, K)
        acc["on"] = False
        L.evx_profile_enable(0)
        launches = int(L.evx_launch_count())
        clocks = sampler.stop()
        fronts = int(L.evx_nds_last_num_fronts())
        if acc["calls"]:
            stages = {k: acc[k] / acc["calls"] for k in ("prep", "tiles", "peel")}
        del ws
        torch.cuda.empty_cache()

        # e2e: host buffers in and out through the C ABI; H2D and D2H inside the timed region
        rank_host = torch.empty(n, dtype=torch.int32).pin_memory()

        def step_e2e():
            check(L.evx_nds_host_f32(x_host.data_ptr(), n, m, rank_host.data_ptr()))
        step_e2e()
        ms_e2e = timed(step_e2e, K)
        assert bool((rank_host.cuda() == rank_out).all()), "host and device entry points disagree"
        verified = verify_last_step(torch, dist, x, rank_out, world, rank, args.verify_rows)
    else:
        from evox_b200.distributed import FusedShardedNonDominatedSort, ShardedNonDominatedSort
        sh = (ShardedNonDominatedSort if args.exchange == "nccl" else FusedShardedNonDominatedSort)(n, m)
        for _ in range(W):
            sh(x)
        sampler = ClockSampler(local)
        sampler.start()
        L.evx_launch_count_reset()
        holder = {}

        def step():
            holder["r"] = sh(x)
        ms = timed(step, K)
        launches = int(L.evx_launch_count())
        clocks = sampler.stop()
        fronts = sh.num_fronts
        rank_host = torch.empty(n, dtype=torch.int32).pin_memory()

        def step_e2e():
            xd = x_host.cuda(non_blocking=True)
            rank_host.copy_(sh(xd), non_blocking=True)
            torch.cuda.current_stream().synchronize()
        step_e2e()
        ms_e2e = timed(step_e2e, K)
        verified = verify_last_step(torch, dist, x, holder["r"], world, rank, args.verify_rows)

    nsga2_multi = None
    if world > 1 and not args.no_nsga2 and not args.no_extras:
        nsga2_multi = nsga2_gens_per_s(torch, dist=dist, exchange=args.exchange)          # every rank takes part
    if not verified["ok"]:
        if rank == 0:
            print(json.dumps({"metric": METRIC, "error": "ranks of the last timed step failed the sampled exactness check",
                              "verified": verified, "n_gpus": world}), flush=True)
        if world > 1:
            dist.destroy_process_group()
        sys.exit(3)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    pairs = float(n) * float(n)
    value = pairs * K / (ms * 1e-3)
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload_name(n, m, args.dist), "seed": args.seed, "fronts": fronts, "parallelism": f"column-panel shards x{world}" + (f", front exchange: {args.exchange}" if world > 1 else ""),
                       "l2": "working set (packed bit matrix %.1f GB) >> 126 MB L2; no flush needed" % (n * n / 16 / 1e9)},
            "clocks": clocks, "gpu_launches": launches, "verified": verified,
            "e2e": {"value": pairs * K / (ms_e2e * 1e-3), "unit": UNIT, "h2d_bytes_per_step": n * m * 4,
                    "d2h_bytes_per_step": n * 4, "ms_per_step": ms_e2e / K}}
    if stages:
        t = stages["tiles"] * 1e-3
        # one bit per ordered pair x the pairs one launch processes: the block-upper triangle, N^2/2
        # (the pre-sort decides the other half without looking at it)
        alg_bytes = pairs / 16.0
        headline = (n == 1_000_000 and m == 3)        # the ncu DRAM counters in profiles/ are for this shape
        line["roofline"] = {"bound": "hbm", "kernel": "k_dominance_tiles", "achieved": alg_bytes / t / 1e9, "peak": hbm_peak,
                            "unit": "GB/s", "frac": alg_bytes / t / 1e9 / hbm_peak,
                            "traffic": None, "traffic_source": None,
                            "reference_equivalent": {"bytes": pairs / 8.0, "GB/s": pairs / 8.0 / t / 1e9,
                                                     "note": "one bit for every one of the reference's N^2 ordered pairs"},
                            "peak_source": peak_src, "ms_per_launch": stages["tiles"],
                            "note": "algorithmic bytes = N^2/16: 1 bit written per tested pair, N^2/2 pairs tested per launch; "
                                    "what limits the kernel is instruction issue, not HBM: see issue_roofline"}
        tp = stages["peel"] * 1e-3
        peel_bytes = pairs / 16.0                     # every stored bit is read exactly once by the peel
        line["peel_roofline"] = {"bound": "hbm", "kernel": "k_peel_dynamic", "achieved": peel_bytes / tp / 1e9,
                                 "peak": hbm_peak, "unit": "GB/s", "frac": peel_bytes / tp / 1e9 / hbm_peak,
                                 "traffic": None, "ms_per_launch": stages["peel"],
                                 "random_gather_peak": GATHER_PEAK_256B,
                                 "frac_of_random_gather_peak": peel_bytes / tp / 1e9 / GATHER_PEAK_256B,
                                 "note": "random gather of 256-byte super-rows through a cp.async ring: "
                                         "tools/probes/gather_probe4.cu measures 6.76 TB/s for exactly this ring and consumer "
                                         "(profiles/r1_gather_probe.txt); the rest is per-front latency (about 230 fronts)"}
        line["stages_ms"] = stages
        ncu = {}
        try:
            ncu = json.loads(NCU_SUMMARY.read_text())
        except Exception:
            pass
        if headline and ncu:
            line["roofline"]["traffic"] = ncu.get("tiles_dram_bytes_per_launch")
            line["roofline"]["traffic_source"] = ncu.get("source")
            line["peel_roofline"]["traffic"] = ncu.get("peel_dram_bytes_per_launch")
        if ncu.get("tiles_issue"):
            # the tile kernel is bound by instruction issue: report what ncu measured for it, not a synthetic peak
            line["issue_roofline"] = dict(ncu["tiles_issue"], kernel="k_dominance_tiles", source=ncu.get("source"),
                                          tested_pairs_per_s=pairs / 2.0 / t, algorithmic_pairs_per_s=pairs / t)
    if nsga2_multi is not None:
        line["nsga2"] = nsga2_multi
    if world == 1 and not args.no_extras:
        if not args.no_nsga2:
            line["nsga2"] = nsga2_gens_per_s(torch)
            line["nsga2_small"] = nsga2_gens_per_s(torch, pop=100, d=7, m=3, gens=100, problem="DTLZ1")   # configs[0] shape
            line["nsga2_mid"] = nsga2_gens_per_s(torch, pop=2500, d=12, m=3, gens=100, problem="DTLZ2")    # merged N = 5000: one-CTA peel
        line["operators"] = operator_rates(torch, hbm_peak)
        line["other_inputs"] = other_inputs(torch, L, n, m, args)
    if not args.no_cpu_baseline:
        ns = auto_cpu_sample(m, args.seed, user=args.cpu_sample)
        _, dt1, _ = cpu_oracle_rate(ns, m, args.seed)
        reps = max(1, min(20, int(12.0 / max(dt1, 1e-3))))             # about 12 s of CPU work in total
        rate, dt, cores = cpu_oracle_rate(ns, m, args.seed, reps=reps)
        line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": cores, "kind": "port",
                                "sample": f"{reps} x non_dominated_sort of {ns} rows ({dt:.2f} s each) with "
                                          f"oracle/oracle_c.c (dense N x N byte matrix + per-front peel, as the reference)"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def operator_rates(torch, hbm_peak):
    """The other rows of the hot path (crowding keys/s, evals/s), inputs resident, CUDA events,
    median of 10 after 3 warm-ups.  GB/s figures use the algorithmic bytes of DESIGN.md section 4."""
    import evox_b200 as E
    g = torch.Generator(device="cuda").manual_seed(1)

    def med(fn, reps=10, inner=10):
        """Median over `reps` of the device time of `inner` back-to-back calls / inner: what the stream
        sustains (a single call between two events mostly measures the host side of the launch)."""
        for _ in range(3):
            fn()
        ts = []
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(inner):
                fn()
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1) / inner)
        ts.sort()
        return ts[len(ts) // 2] * 1e-3

    out = {}
    n, m = 200_000, 3
    F = torch.rand((n, m), device="cuda", generator=g)
    t = med(lambda: E.operators.crowding_distance(F))
    out["crowding_distance"] = {"workload": f"N={n} M={m} mask=None", "keys_per_s": n * m / t, "ms": t * 1e3,
                                "GB/s": (4 * n * m + 4 * n) / t / 1e9}
    mask = torch.rand(n, device="cuda", generator=g) < 0.02
    t = med(lambda: E.operators.crowding_distance(F, mask))
    out["crowding_distance_masked"] = {"workload": f"N={n} M={m} mask = 2% (a cut front)", "keys_per_s": n * m / t,
                                       "ms": t * 1e3}
    x = torch.rand((16384, 3), device="cuda", generator=g)
    t = med(lambda: E.utils.dominate_relation(x, x))
    out["dominate_relation"] = {"workload": "16384 x 16384, M=3, bool matrix out", "pairs_per_s": 16384.0 ** 2 / t,
                                "GB/s": 16384.0 ** 2 / t / 1e9, "frac_of_hbm_peak": 16384.0 ** 2 / t / 1e9 / hbm_peak, "ms": t * 1e3}
    for name, nn, d, mm in (("DTLZ2", 100_000, 12, 3), ("MaF1", 50_000, 19, 10), ("LSMOP1", 100_000, 300, 3)):
        X = torch.rand((nn, d), device="cuda", generator=g)
        prob = getattr(E.problems.numerical, name)(d=d, m=mm)
        t = med(lambda: prob.evaluate(None, X))
        b = 4.0 * nn * (d + mm)
        out[f"evaluate_{name}"] = {"workload": f"n={nn} d={d} m={mm}", "evals_per_s": nn / t, "GB/s": b / t / 1e9,
                                   "frac_of_hbm_peak": b / t / 1e9 / hbm_peak, "us": t * 1e6}
    # NSGA-III environmental selection at the configs[3] shape (MaF1-like: merged n = 100k, M = 10,
    # 48 620 reference directions = Das-Dennis H = 9)
    from evox_b200.algorithms import nsga3, spea2
    n, m, R, ps = 100_000, 10, 48_620, 50_000
    fit = torch.rand((n, m), device="cuda", generator=g)
    ref = torch.rand((R, m), device="cuda", generator=g) + 0.01
    rank = E.operators.non_dominated_sort(fit)
    last_rank = int(rank[torch.sort(rank, stable=True).indices[ps]])
    norm = nsga3.normalize(fit, rank, last_rank)
    gid, gd = nsga3.associate(norm, ref)
    rw = torch.randint(0, 2**31 - 1, (ps + 1, 2), device="cuda", dtype=torch.int32, generator=g)
    pop = torch.rand((n, 19), device="cuda", generator=g)
    ta = med(lambda: nsga3.associate(norm, ref), reps=5, inner=2)
    tn = med(lambda: nsga3.niche(rank, last_rank, gid, gd, R, ps, rw), reps=5, inner=2)
    te = med(lambda: nsga3.environmental_selection(pop, fit, ref, ps, rw), reps=5, inner=1)
    out["nsga3_selection"] = {"workload": f"merged n={n} M={m} refs={R} pop={ps}", "associate_ms": ta * 1e3,
                              "associate_pairs_per_s": n * R / ta, "niche_ms": tn * 1e3,
                              "niche_picks": ps - int((rank < last_rank).sum()), "whole_selection_ms": te * 1e3}
    # SPEA2 fitness assignment and archive truncation (merged n = 20k, M = 3, 60 % valid cut to 50 %)
    n, m = 20_000, 3
    obj = torch.rand((n, m), device="cuda", generator=g)
    mask = torch.rand(n, device="cuda", generator=g) < 0.6
    k = int(mask.sum()) - n // 2
    tf = med(lambda: spea2.cal_fitness(obj), reps=5, inner=2)
    tt = med(lambda: spea2._truncation(obj, k, mask), reps=3, inner=1)
    out["spea2"] = {"workload": f"n={n} M={m}, truncation removes {k} rows", "cal_fitness_ms": tf * 1e3,
                    "cal_fitness_pairs_per_s": float(n) * n / tf, "truncation_ms": tt * 1e3}
    return out


def other_inputs(torch, L, n, m, args):
    """the same sort on seeds {0, 1, 2} and on the structured inputs that drive the peel
    differently (one front = a converged NSGA-II population), ms per sort, 3 timed calls after 2 warm-ups."""
    from evox_b200._lib import check
    out = {}
    ws = torch.empty(L.evx_nds_workspace_bytes(n, m), dtype=torch.uint8, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    cases = [("uniform", sd, n) for sd in (0, 1, 2)] + [("one_front", 0, n)]
    if n != 200_000:
        cases += [("uniform", 0, 200_000), ("one_front", 0, 200_000)]
    for dist, seed, nn in cases:
        x = torch.from_numpy(make_input(nn, m, seed, dist)).cuda()
        r = torch.empty(nn, dtype=torch.int32, device="cuda")

        def call():
            check(L.evx_nds_f32(x.data_ptr(), nn, m, r.data_ptr(), ws.data_ptr(), ws.numel(), s))
        for _ in range(2):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(3):
            call()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        out[f"{dist}_seed{seed}_n{nn}"] = {"ms": ms, "pairs_per_s": float(nn) ** 2 / (ms * 1e-3), "fronts": int(r.max()) + 1}
    return out


def nsga2_gens_per_s(torch, pop=100_000, d=12, m=3, gens=50, problem="DTLZ2", dist=None, exchange="fused"):
    """Second half of BASELINE.json's metric: NSGA-II generations/s at pop 100k on DTLZ2 (configs[2]):
    variation kernel -> DTLZ2 evaluate -> concat -> survivor selection (merged N = 2 pop -> pop), all on the
    device, >= 50 generations after 3 of warm-up.  With several GPUs (dist given): state
    replicated, evaluation sharded by rows + all-gather of the fitness, the sort sharded by column panels
    (evox_b200.distributed.ShardedNSGA2); timed between barriers, max over ranks."""
    import evox_b200 as E
    from evox_b200.algorithms import NSGA2, run_generations
    alg = NSGA2(torch.zeros(d), torch.ones(d), n_objs=m, pop_size=pop)
    prob = getattr(E.problems.numerical, problem)(d=d, m=m)
    world = dist.get_world_size() if dist is not None else 1
    if world > 1:
        from evox_b200.distributed import ShardedNSGA2
        drv = ShardedNSGA2(alg, exchange=exchange)
        run = lambda st, g: drv.run(prob, st, g)[0]
    else:
        run = lambda st, g: run_generations(alg, prob, st, g)[0]
    state = run(alg.setup(seed=0), 3)                                    # warm-up (first fronts are the deepest)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    e0.record()
    state = run(state, gens)
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    chk = torch.tensor([float(state.fitness.double().sum())], device="cuda", dtype=torch.float64)
    same = True
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        lo, hi = chk.clone(), chk.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        same = bool((lo == hi).all())
    ms = float(ms)
    return {"metric": "nsga2_generations_per_s", "value": gens / (ms * 1e-3), "unit": "generations/s", "n_gpus": world,
            "ms_per_generation": ms / gens, "config": {"workload": f"NSGA-II {problem} d={d} M={m} pop={pop} (merged N={2 * pop})",
                                                       "generations_timed": gens, "after_warmup_generations": 3},
            "population_identical_on_all_gpus": same, "fitness_checksum": float(chk),
            "note": "variation = evx_variation_sbx_pm_f32 (Philox; value parity with the reference's operators for a given "
                    "draw), selection = evx_non_dominate_f32 / evx_select_by_rank_f32 (non_dominate.py:210-222 as one call)"}


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
