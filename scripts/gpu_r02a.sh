# round-2 run A (one GPU): parity suite, window step at N=1 and the per-rank compute of an N=8 step emulated on this GPU
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
for s in 1 2; do
python scripts/bench_window.py --steps 10 --warmup 3 --streams $s 2>>gpurun_out/a.err | tail -1 | cut -c1-300
python scripts/bench_window.py --steps 10 --warmup 3 --streams $s --emulate 8 2>>gpurun_out/a.err | tail -1 | cut -c1-400
done
python scripts/bench_window.py --steps 10 --warmup 3 --streams 1 --emulate 8 --plan strips 2>>gpurun_out/a.err | tail -1 | cut -c1-400
python bench.py --workload config3 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>>gpurun_out/a.err | tail -1 > gpurun_out/c3_view.json; python -c "
import json; d=json.load(open('gpurun_out/c3_view.json')); print(d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()}, d['counts'])"
tail -5 gpurun_out/a.err
python bench.py --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_ours.json 2>>gpurun_out/a.err; cut -c1-3000 gpurun_out/bench_ours.json
tail -5 gpurun_out/a.err
