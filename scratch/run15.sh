python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_v11.json 2> gpurun_out/nerf_v11.err || tail -5 gpurun_out/nerf_v11.err
python bench.py --workload sdf --no-cpu-baseline > gpurun_out/sdf_v4.json 2> gpurun_out/sdf_v4.err || tail -5 gpurun_out/sdf_v4.err
python bench.py --workload render --no-cpu-baseline > gpurun_out/render_v5.json 2> gpurun_out/render_v5.err || tail -5 gpurun_out/render_v5.err
