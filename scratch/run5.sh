python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_v8.json 2> gpurun_out/nerf_v8.err || tail -5 gpurun_out/nerf_v8.err
python bench.py --workload hashgrid --no-cpu-baseline > gpurun_out/bench_hashgrid_v3.json 2> gpurun_out/bench_hashgrid_v3.err
