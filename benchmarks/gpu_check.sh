# scratch helper for one gpurun call (rewritten per run): gradient all-reduce on a high-priority NCCL stream, A/B
set -x
N=${1:-2}
mkdir -p gpurun_out
run() { # name, env, flags
  timeout 150 env $2 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29530 + RANDOM % 400)) bench.py --gpus $N $3 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/$1.json 2> gpurun_out/$1.err
  tail -1 gpurun_out/$1.err | cut -c1-300; cut -c1-160 gpurun_out/$1.json
}
run sdf_n${N}_p0 NGP_GRADIENT_GROUP_PRIORITY=0 "--workload sdf --only"
run sdf_n${N}_p1 NGP_GRADIENT_GROUP_PRIORITY=1 "--workload sdf --only"
run sdf_n${N}_p1c2 NGP_GRADIENT_GROUP_PRIORITY=1 "--workload sdf --only --sdf-chunks 2"
run sdf_n${N}_p1c4 NGP_GRADIENT_GROUP_PRIORITY=1 "--workload sdf --only --sdf-chunks 4"
run nerf_n${N}_p0 NGP_GRADIENT_GROUP_PRIORITY=0 ""
run nerf_n${N}_p1 NGP_GRADIENT_GROUP_PRIORITY=1 ""
