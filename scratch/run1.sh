set -x
python -m pytest tests/test_gpu_hashgrid.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -5
python benchmarks/hashgrid_micro.py > gpurun_out/hg_micro_rand.txt 2>&1
python benchmarks/hashgrid_micro.py --coherent --B 589824 > gpurun_out/hg_micro_coh.txt 2>&1
python bench.py --workload hashgrid --no-cpu-baseline > gpurun_out/bench_hashgrid_v3.json 2> gpurun_out/bench_hashgrid_v3.err
python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/bench_nerf_v7.json 2> gpurun_out/bench_nerf_v7.err
tail -2 gpurun_out/bench_nerf_v7.err
