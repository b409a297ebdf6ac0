set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_nerf_step.py tests/test_abi.py -x -q > gpurun_out/pytest_a.log 2>&1; tail -15 gpurun_out/pytest_a.log
timeout 200 python bench.py --only --no-cpu-baseline > gpurun_out/bench_train.json 2> gpurun_out/bench_train.err; tail -3 gpurun_out/bench_train.err
timeout 200 python bench.py --only --no-cpu-baseline --workload render --steps 8 > gpurun_out/bench_render.json 2> gpurun_out/bench_render.err; tail -3 gpurun_out/bench_render.err
