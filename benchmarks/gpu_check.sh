# one GPU-box pass over the changed pieces (development helper; bench.py and tests/ are the contract)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ffmlp.py tests/test_reference_wrappers.py tests/test_vs_reference_ext.py tests/test_golden.py -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_a.log
tail -25 gpurun_out/pytest_a.log
