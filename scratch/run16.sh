python -m pytest tests/test_gpu_raymarching.py tests/test_gpu_nerf_step.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_v12.json 2> gpurun_out/nerf_v12.err || tail -5 gpurun_out/nerf_v12.err
