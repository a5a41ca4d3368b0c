"""ICP throughput probe (BASELINE config 4): python tools/icp_probe.py [pairs] [max_dist]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import bench_extra
n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
r = float(sys.argv[2]) if len(sys.argv) > 2 else 0.004
out = bench_extra.icp_bench(n, r)
out["pairs_per_s"] = n / (out["ms"] * 1e-3)
print(json.dumps(out))
