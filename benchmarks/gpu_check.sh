set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; tail -3 gpurun_out/pytest_a.log
( time timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2> gpurun_out/bench_default.time; tail -3 gpurun_out/bench_default.err; cat gpurun_out/bench_default.time
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -2 gpurun_out/bench_reference.err
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
