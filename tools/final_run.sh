#!/bin/bash
# usage (on a one-GPU box): tools/final_run.sh  -> full GPU suite, smoke, the bench matrix, then the ncu launch list and
# --set full captures of the tile and peel kernels of the same bench command (after its plain run exited 0)
O=gpurun_out
python -m pytest tests -m gpu -x -q > $O/r2_final_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r2_final_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > $O/r2_bench_n1m.json 2> $O/r2_bench_n1m.err; echo "bench rc=$?"
python bench.py --impl reference > $O/r2_bench_reference.json 2>> $O/r2_bench_n1m.err
for n in 200000 100000 10000; do python bench.py --rows $n --steps 20 --no-extras --no-nsga2 > $O/r2_bench_n${n}_m3.json 2>> $O/r2_bench_n1m.err; done
python bench.py --rows 100000 --m 10 --steps 20 --no-extras --no-nsga2 > $O/r2_bench_n100000_m10.json 2>> $O/r2_bench_n1m.err
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras --no-nsga2"
$B > $O/plain.log 2>&1 && {
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/r2_launches_n1m.csv $B > $O/ncu_l.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:k_dominance_tiles -s 3 -c 1 -f -o $O/r2_tiles_n1m $B > $O/ncu_t.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:k_peel_dynamic -s 3 -c 1 -f -o $O/r2_peel_n1m $B > $O/ncu_p.log 2>&1
}
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f, d.get('ms_per_step'), 'e2e', (d.get('e2e') or {}).get('ms_per_step'), 'frac', (d.get('roofline') or {}).get('frac'), 'verified', (d.get('verified') or {}).get('ok'), 'nsga2', (d.get('nsga2') or {}).get('value'))
    except Exception as e: print(f,'FAILED',e)
PY
ls -la $O/*.ncu-rep
