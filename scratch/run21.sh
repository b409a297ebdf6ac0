for o in 0 8192 32768 131072 524288; do
python bench.py --workload render --no-cpu-baseline --opt march_rays_warp=$o > gpurun_out/render_w$o.json 2>/dev/null
done
