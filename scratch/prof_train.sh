# profiles of the training step (configs[1]); run under gpurun from the repo root
set -x
python bench.py --steps 4 --warmup 3 --pretrain 48 --no-cpu-baseline --no-graph --no-overlap > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r1_nerf_train_launches_v2.csv \
    python bench.py --steps 4 --warmup 3 --pretrain 48 --no-cpu-baseline --no-graph --no-overlap > gpurun_out/prof_launches.log 2>&1
python scratch/trainer_small.py || exit 1
ncu --set full --clock-control none --import-source on -k regex:'k_ffmlp_bwd_mp|k_grid_bwd_v3|k_grid_fwd_v3|k_adam|k_march_train_count_warp|k_ffmlp_fwd' --launch-skip 16 -c 8 \
    -o gpurun_out/prof_nerf_train_v2 -f python scratch/trainer_small.py > gpurun_out/prof_full.log 2>&1
tail -3 gpurun_out/prof_full.log
