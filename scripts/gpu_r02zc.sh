# round-2 run ZC (one GPU): single-pass preprocess_bwd variants -- suite, stage times at configs 2 / 3 / 4
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5
for w in config3 config4; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/zc.err | tail -1 > gpurun_out/zc_$w.json
  python -c "
import json; d=json.load(open('gpurun_out/zc_$w.json')); print('$w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
