set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_nerf_step.py -x -q > gpurun_out/pytest_a.log 2>&1; tail -12 gpurun_out/pytest_a.log
for o in 1 0 1 0; do timeout 200 python bench.py --only --no-cpu-baseline --opt nerf_nets_tmem=$o 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); ks=d['kernels']; k=[v for n,v in ks.items() if n.startswith('k_nerf_nets_fwd')][0]; s=d['config']['samples_per_step_per_gpu']
print('TMEM=$o ms/step', round(d['ms_per_step'],4), 'nets_fwd us', round(k['us_per_launch'],1), 'samples', s, 'ps/sample', round(k['us_per_launch']*1e6/s,1))"; done
for o in 1 0; do timeout 200 python bench.py --only --no-cpu-baseline --workload render --steps 6 --opt nerf_nets_tmem=$o 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); ks=d['kernels']; k=[v for n,v in ks.items() if n.startswith('k_nerf_nets_fwd')][0]
print('TMEM=$o render ms/frame', round(d['ms_per_step'],2), 'nets_fwd us', round(k['us_per_launch'],1))"; done
