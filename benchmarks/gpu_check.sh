set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nerf_step.py tests/test_gpu_sdf_step.py tests/test_gpu_concurrency.py -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; tail -15 gpurun_out/pytest_a.log
timeout 200 python bench.py --only --no-cpu-baseline > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; tail -3 gpurun_out/bench_train.err
timeout 200 python bench.py --only --no-cpu-baseline --workload sdf > gpurun_out/bench_sdf.json 2> gpurun_out/bench_sdf.err; tail -3 gpurun_out/bench_sdf.err
