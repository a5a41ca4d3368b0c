#!/bin/bash
# Run bench.py once per tuning variant in lib/variants/ (on the GPU box); prints kernel ms per launch.
cd "$(dirname "$0")/.."
for so in 3d-reconstruction-planning_b200/lib/variants/lib_*.so; do
  for wl in tsdf1024 tsdf512; do
    B2R_LIB=$PWD/$so timeout 300 python bench.py --steps 32 --warmup 3 --no-cpu-baseline --workload $wl 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$so'.split('/')[-1], '$wl', 'kernel_ms', round(d['roofline']['kernel_ms_per_launch'],4), 'frac', round(d['roofline']['frac'],3), 'fps', round(d['value'],1), 'e2e', round(d['e2e']['value'],1))"
  done
done
