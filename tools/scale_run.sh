#!/bin/bash
# usage: tools/scale_run.sh N   (on a box with N GPUs): correctness check, then bench at N = 1M and 200k
N=$1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 400 $T tools/fused_check.py > gpurun_out/r2_fused_check_world$N.log 2>&1; grep -v "Warning\|\*\*\*\|OMP" gpurun_out/r2_fused_check_world$N.log | tail -6
timeout 300 $T bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_scale_g$N.json 2> gpurun_out/r2_scale_g$N.err
timeout 300 $T bench.py --gpus $N --rows 200000 --steps 20 --warmup 3 --no-cpu-baseline --no-nsga2 > gpurun_out/r2_scale_g${N}_200k.json 2>> gpurun_out/r2_scale_g$N.err
timeout 300 $T tools/fused_breakdown.py 1000000 2>&1 | grep -v "Warning\|\*\*\*\|OMP\|NCCL\|^$" > gpurun_out/r2_breakdown_g$N.txt
python - <<PY
import json
for f in ("gpurun_out/r2_scale_g$N.json", "gpurun_out/r2_scale_g${N}_200k.json"):
    try:
        d = json.loads([l for l in open(f).read().splitlines() if l.startswith("{")][-1])
        print(f, d["n_gpus"], "ms", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "verified", d["verified"]["ok"], "nsga2", d.get("nsga2", {}).get("value"), d.get("nsga2", {}).get("population_identical_on_all_gpus"))
    except Exception as e:
        print(f, "FAILED", e)
PY
cat gpurun_out/r2_breakdown_g$N.txt
