python bench.py --workload sdf --no-cpu-baseline --opt hashgrid_bwd_variant=1 > gpurun_out/sdf_bv1.json 2>gpurun_out/sdf_bv1.err || tail -3 gpurun_out/sdf_bv1.err
python bench.py --workload sdf --no-cpu-baseline --opt hashgrid_bwd_variant=1 --opt hashgrid_smem_levels=0 > gpurun_out/sdf_bv1_nosmem.json 2>/dev/null
