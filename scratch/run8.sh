timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python bench.py --workload sdf --no-cpu-baseline > gpurun_out/sdf_v2.json 2> gpurun_out/sdf_v2.err || tail -5 gpurun_out/sdf_v2.err
python bench.py --workload render --no-cpu-baseline > gpurun_out/render_v2.json 2> gpurun_out/render_v2.err || tail -5 gpurun_out/render_v2.err
