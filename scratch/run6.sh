python -m pytest tests/test_gpu_nerf_step.py -x -q -m gpu 2>&1 | tail -4
for ab in 2 3 5 0; do
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --adam-blocks $ab > gpurun_out/nerf_v9_ab$ab.json 2> gpurun_out/nerf_v9_ab$ab.err || tail -5 gpurun_out/nerf_v9_ab$ab.err
done
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-overlap > gpurun_out/nerf_v9_seq.json 2> gpurun_out/nerf_v9_seq.err || tail -5 gpurun_out/nerf_v9_seq.err
