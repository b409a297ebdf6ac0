for r in 0 2; do
python bench.py --workload sdf --no-cpu-baseline --opt hashgrid_bwd_red4=$r > gpurun_out/sdf_red4_$r.json 2>/dev/null
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --opt hashgrid_bwd_red4=$r > gpurun_out/nerf_red4_$r.json 2>/dev/null
done
