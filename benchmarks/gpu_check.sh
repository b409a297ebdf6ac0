set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ffmlp.py tests/test_gpu_nerf_step.py tests/test_gpu_sdf_step.py tests/test_reference_wrappers.py tests/test_golden.py -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; tail -12 gpurun_out/pytest_a.log
for o in 1 0 1 0; do timeout 200 python bench.py --only --no-cpu-baseline --workload sdf --opt ffmlp_fwd_tmem=$o 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); ks=d['kernels']; k=[v for n,v in ks.items() if n.startswith('k_ffmlp_fwd')][0]
print('TMEM=$o sdf ms/step', round(d['ms_per_step'],4), 'ffmlp_fwd us', round(k['us_per_launch'],1))"; done
