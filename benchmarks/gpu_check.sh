# scratch helper for one gpurun call (rewritten per run)
set -x
mkdir -p gpurun_out
for i in 1 2 3; do
  timeout 200 python bench.py --only --no-cpu-baseline 2>/dev/null > gpurun_out/nt_$i.json; python -c "
import json;d=json.load(open('gpurun_out/nt_$i.json'));print(d['ms_per_step'], d['step_ms']['median'], d['step_ms']['min'], d['step_ms']['max'])"
done
