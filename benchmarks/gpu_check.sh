set -x
mkdir -p gpurun_out
for u in 1 0 1 0; do timeout 200 python bench.py --only --no-cpu-baseline --opt hashgrid_v3_unroll_bwd=$u 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); k=d['kernels']['k_grid_bwd_v3']; s=d['config']['samples_per_step_per_gpu']
print('unroll=$u ms/step', round(d['ms_per_step'],4), 'grid_bwd us', round(k['us_per_launch'],1), 'samples', s, 'ps/sample', round(k['us_per_launch']*1e6/s,1))"; done
