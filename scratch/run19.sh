python -m pytest tests/test_gpu_ffmlp.py tests/test_vs_reference_ext.py -q -m gpu 2>&1 | grep -v "^$" | tail -14 | cut -c1-220
python scratch/ffmlp_bwd_sweep.py 2>&1 | grep "variant 1 split 2 issuers\|variant 1 split 1 issuers 2" 
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_v13.json 2>/dev/null
