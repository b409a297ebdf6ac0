# one GPU-box pass over the changed pieces (development helper; bench.py and tests/ are the contract)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_nerf_step.py tests/test_gpu_ffmlp.py tests/test_gpu_sdf_step.py tests/test_gpu_concurrency.py -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_a.log
tail -25 gpurun_out/pytest_a.log
python bench.py --only --no-cpu-baseline > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; tail -3 gpurun_out/bench_train.err
python bench.py --only --no-cpu-baseline --workload render --steps 8 > gpurun_out/bench_render.json 2> gpurun_out/bench_render.err; tail -3 gpurun_out/bench_render.err
