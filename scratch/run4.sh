python -m pytest tests/test_gpu_hashgrid.py -x -q -m gpu 2>&1 | tail -2
for o in "hashgrid_bwd_red4=0" "hashgrid_v3_mix_bwd=0" "hashgrid_v3_minb_bwd=6" "hashgrid_v3_minb_fwd=4"; do
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --opt $o > gpurun_out/nerf_$o.json 2> gpurun_out/nerf_$o.err || tail -5 gpurun_out/nerf_$o.err
done
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --opt hashgrid_v3_minb_bwd=6 --opt hashgrid_v3_mix_bwd=0 > gpurun_out/nerf_hashgrid_b6m0.json 2>/dev/null
