python -m pytest tests/test_gpu_hashgrid.py -x -q -m gpu 2>&1 | tail -3
for o in 0 1; do
python bench.py --workload sdf --no-cpu-baseline --opt hashgrid_bwd_smem_coarse=$o > gpurun_out/sdf_sc$o.json 2>/dev/null
python bench.py --workload hashgrid --no-cpu-baseline --opt hashgrid_bwd_smem_coarse=$o > gpurun_out/hg_sc$o.json 2>/dev/null
done
