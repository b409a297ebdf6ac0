# one GPU-box pass over the changed pieces (development helper; bench.py and tests/ are the contract)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_a.log
tail -6 gpurun_out/pytest_a.log
