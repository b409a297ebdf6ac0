# scratch helper for one gpurun call (rewritten per run): memcheck of the binning kernels and the level-range backward
set -x
mkdir -p gpurun_out
timeout 400 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_gpu_sdf_step.py -x -q -k "bin_points or level_ranges or binned_step" > gpurun_out/sanitizer_sdf.log 2>&1
echo "rc=$?"; tail -6 gpurun_out/sanitizer_sdf.log
