# round-2 run N (one GPU): suite after the pre_bwd variant split and the configurable peer launch shape; stage times
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5
for w in config2 config3; do
  python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/n.err | tail -1 > gpurun_out/n_$w.json
  python -c "
import json; d=json.load(open('gpurun_out/n_$w.json')); print('$w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
python scripts/strip_stages.py --steps 5 --only 0,7,atomic | tail -1
python scripts/strip_stages.py --steps 5 --only 0,30,atomic | tail -1
tail -3 gpurun_out/n.err
