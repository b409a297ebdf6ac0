python bench.py --workload hashgrid --no-cpu-baseline > gpurun_out/prof_hg_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/r1_hashgrid_launches_v3.csv python bench.py --workload hashgrid --no-cpu-baseline > gpurun_out/prof_hg_launches.log 2>&1
