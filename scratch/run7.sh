timeout 300 python -m pytest tests/test_gpu_ffmlp.py tests/test_gpu_umma.py -x -q -m gpu 2>&1 | tail -4
timeout 120 python scratch/ffmlp_ablate.py
