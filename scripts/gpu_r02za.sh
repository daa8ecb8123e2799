# round-2 run ZA (one GPU): sort kernels with medium-bucket lists -- suite, config 5 as stated at N=1, grown-scene launch list
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -5
timeout 900 python scripts/bench_config5.py --reps 3 > gpurun_out/c5_n1za.json 2> gpurun_out/c5_n1za.err; tail -2 gpurun_out/c5_n1za.err; python -c "
import json; d=json.loads(open('gpurun_out/c5_n1za.json').read().strip().splitlines()[-1]); print(list(zip(d['rows'], d['all_steps_ms']))[::5], d['all_steps_ms'][-1], sum(d['all_steps_ms']))"
for w in config2 config3; do
  python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/za.err | tail -1 > gpurun_out/za_$w.json
  python -c "
import json; d=json.load(open('gpurun_out/za_$w.json')); print('$w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02za_c5_grown_1m_launches.csv python scripts/bench_config5.py --stop-at 31 --reps 1 --eager --streams 1 --profile-last > gpurun_out/ncu_za.log 2>&1
grep -c tile_sort_small gpurun_out/r02za_c5_grown_1m_launches.csv
