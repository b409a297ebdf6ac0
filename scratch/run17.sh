for i in 1 2; do
python bench.py --steps 200 --warmup 10 --no-cpu-baseline > gpurun_out/nerf_fl$i.json 2>/dev/null
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-flush > gpurun_out/nerf_nofl$i.json 2>/dev/null
done
