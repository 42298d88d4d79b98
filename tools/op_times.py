"""Device times of the secondary operators (same measurements as bench.py's `operators` field)."""
import json, sys
sys.path.insert(0, '.')
import torch
import bench
peaks = json.load(open('MEASURED_PEAKS.json')) if __import__('os').path.exists('MEASURED_PEAKS.json') else {}
print(json.dumps(bench.operator_rates(torch, float(peaks.get('hbm_gbs', 6650.0))), indent=1))
