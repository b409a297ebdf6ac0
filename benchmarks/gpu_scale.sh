# N-GPU scaling run of bench.py (development helper): bash benchmarks/gpu_scale.sh <N>
set -x
N=${1:-8}
mkdir -p gpurun_out
timeout ${TMO:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -3 gpurun_out/bench_n$N.err
cat gpurun_out/bench_n$N.json | cut -c1-600
