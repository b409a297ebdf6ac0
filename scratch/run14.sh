python -m pytest tests/test_gpu_nerf_step.py -x -q -m gpu 2>&1 | tail -4
python bench.py --workload render --no-cpu-baseline > gpurun_out/render_v4.json 2> gpurun_out/render_v4.err || tail -5 gpurun_out/render_v4.err
python bench.py --workload render --no-cpu-baseline --points 678 > gpurun_out/render_v4_small.json 2> gpurun_out/render_v4_small.err || tail -5 gpurun_out/render_v4_small.err
