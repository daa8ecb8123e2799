# quick iteration: parity tests + own-arm bench (no CPU baseline)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 100 --warmup 10 --no-cpu-baseline $BENCH_ARGS > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; tail -3 gpurun_out/bench_quick.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_quick.json'))
print('value',d['value'],'ms/step',d['ms_per_step'],'e2e',d.get('e2e',{}).get('value'), 'clk', d['clocks'])
for k,v in d['stages'].items(): print(f"  {k:16s} {v['ms']*1e3:8.1f} us  {v['gbs']:8.1f} GB/s")
PY
