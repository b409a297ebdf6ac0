python -m pytest tests/test_gpu_ffmlp.py tests/test_gpu_nerf_step.py tests/test_gpu_sdf_step.py tests/test_vs_reference_ext.py -x -q -m gpu 2>&1 | tail -2
timeout 200 python scratch/ffmlp_split_sweep.py | grep "issuers 3 split 0"
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_v14.json 2>/dev/null
