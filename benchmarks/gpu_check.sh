set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_a.log 2>&1; tail -4 gpurun_out/pytest_a.log
( time timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err ) 2> gpurun_out/bench_default.time; tail -3 gpurun_out/bench_default.err; cat gpurun_out/bench_default.time
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python benchmarks/step_prof.py 1 > gpurun_out/plain2.log 2>&1 && ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"k_grid_fwd_v3|k_grid_bwd_v3|k_ffmlp|k_nerf_nets|k_march_train_count|k_adam" -o gpurun_out/r2_step_full_v2 python benchmarks/step_prof.py 1 > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/ncu2.log
python benchmarks/step_prof.py 2 > gpurun_out/plain.log 2>&1 && ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_step_launches.csv python benchmarks/step_prof.py 2 > gpurun_out/ncu1.log 2>&1
