set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ffmlp.py tests/test_gpu_nerf_step.py tests/test_gpu_sdf_step.py -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; tail -12 gpurun_out/pytest_a.log
for o in 1 0 2 1 0 2; do timeout 200 python bench.py --only --no-cpu-baseline --opt ffmlp_bwd_tmem=$o 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); ks=d['kernels']; s=d['config']['samples_per_step_per_gpu']
a=ks['k_ffmlp_bwd_mp_color']['us_per_launch']; b=ks['k_ffmlp_bwd_mp']['us_per_launch']
print('BTM=$o ms/step', round(d['ms_per_step'],4), 'bwd colour us', round(a,1), 'sigma us', round(b,1), 'samples', s, 'ps/sample', round(a*1e6/s,1), round(b*1e6/s,1))"; done
