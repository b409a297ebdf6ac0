python -m pytest tests/test_gpu_hashgrid.py -x -q -m gpu 2>&1 | tail -3
python benchmarks/hashgrid_micro.py > gpurun_out/hg_micro_rand.txt 2>&1
python benchmarks/hashgrid_micro.py --coherent --B 589824 > gpurun_out/hg_micro_coh.txt 2>&1
