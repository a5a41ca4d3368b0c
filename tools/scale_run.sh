N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo rc=$?
python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_n$N.json').read().strip().splitlines()[-1])
print('N=$N fps', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['config'].get('slab_boundaries'), d.get('rank_ms_per_step')); m=d['mesh']; print({k:round(m[k],3) for k in m if k.endswith('_ms')}, m['checksum'], m.get('per_slab_vertices')); print(d['icp']['pairs_per_s'], d['icp_r0.05']['pairs_per_s']); print(d.get('tsdf2048'))"
