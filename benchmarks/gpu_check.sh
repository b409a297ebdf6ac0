# scratch helper for one gpurun call (rewritten per run): SDF step at N GPUs, final defaults
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --workload sdf --only --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/sdf_final_n$N.json 2> gpurun_out/sdf_final_n$N.err
tail -1 gpurun_out/sdf_final_n$N.err | cut -c1-300; cut -c1-200 gpurun_out/sdf_final_n$N.json
