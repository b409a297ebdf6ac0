python scratch/hg_prof.py || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_grid -o gpurun_out/hg_v3 -f python scratch/hg_prof.py > gpurun_out/hg_v3_ncu.log 2>&1
tail -3 gpurun_out/hg_v3_ncu.log
