"""ScalableTSDFVolume throughput probe (row f4): python tools/scalable_probe.py [resolution] [frames]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools"), os.path.join(ROOT, "3d-reconstruction-planning_b200")]
import bench_extra
R = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
nf = int(sys.argv[2]) if len(sys.argv) > 2 else 64
steps = int(sys.argv[3]) if len(sys.argv) > 3 else None
print(json.dumps(bench_extra.scalable_bench(R, nf, 0.9 if R >= 1024 else 0.5, steps)))
