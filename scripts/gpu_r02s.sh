# round-2 run S (one GPU): bucket sort in the large-tile kernel -- parity suite, config 5 at its final size, configs 3 / 4
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5
python scripts/bench_config5.py --fixed-rows 2000000 --reps 5 2>gpurun_out/s.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c5 2M n1', d['all_steps_ms'])"
python scripts/bench_config5.py --fixed-rows 2000000 --reps 5 --emulate 8 2>>gpurun_out/s.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c5 2M rank0 of 8', d['all_steps_ms'])"
python scripts/bench_config5.py --fixed-rows 1000000 --reps 5 2>>gpurun_out/s.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('c5 1M n1', d['all_steps_ms'])"
for w in config3 config4; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/s.err | tail -1 > gpurun_out/s_$w.json
  python -c "
import json; d=json.load(open('gpurun_out/s_$w.json')); print('$w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
python scripts/bench_window.py --workload config5 --steps 5 --warmup 2 2>>gpurun_out/s.err | tail -1 | cut -c1-200
tail -3 gpurun_out/s.err
