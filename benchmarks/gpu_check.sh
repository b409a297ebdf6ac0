# one GPU-box pass over the changed pieces (development helper; bench.py and tests/ are the contract)
set -x
mkdir -p gpurun_out
( time python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2> gpurun_out/bench_default.time
tail -3 gpurun_out/bench_default.err
python benchmarks/ffmlp_micro.py > gpurun_out/ffmlp_micro.log 2>&1
cat gpurun_out/ffmlp_micro.log | cut -c1-400
