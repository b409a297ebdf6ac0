# scratch helper for one gpurun call (rewritten per run): SDF step at N GPUs, pipelined table-gradient all-reduce A/B
set -x
N=${1:-2}
mkdir -p gpurun_out
timeout 120 python bench.py --workload sdf --only --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/sdf_n1.json 2>gpurun_out/sdf_n1.err; cut -c1-200 gpurun_out/sdf_n1.json
for c in 1 4 2; do
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$c bench.py --workload sdf --only --gpus $N --sdf-chunks $c --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/sdf_n${N}_c$c.json 2> gpurun_out/sdf_n${N}_c$c.err
  tail -2 gpurun_out/sdf_n${N}_c$c.err; cut -c1-200 gpurun_out/sdf_n${N}_c$c.json
done
