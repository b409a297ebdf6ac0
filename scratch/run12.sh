( time python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2> gpurun_out/bench_default.time
tail -3 gpurun_out/bench_default.time
( time python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2> gpurun_out/bench_ref.time
tail -3 gpurun_out/bench_ref.time
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
nproc
