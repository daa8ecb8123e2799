# round-2 run B (one GPU): parity suite incl. arena tests, config 5 as stated at N=1
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 900 python scripts/bench_config5.py --reps 3 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; tail -5 gpurun_out/c5_n1.err; cut -c1-1800 gpurun_out/c5_n1.json
# kernel A/B: compositing-kernel variants (scripts/variants), parity tests then stage times at configs 2 and 3
for v in base fwdc bwdd both; do
  if [ $v = base ]; then unset WGS_RASTERIZER_LIB; else export WGS_RASTERIZER_LIB=$PWD/scripts/variants/libwgs_$v.so; fi
  if [ $v != base ]; then python -m pytest tests/test_gpu_render.py tests/test_gpu_api.py tests/test_gpu_backward.py -m gpu -x -q 2>&1 | tail -3; fi
  for w in config2 config3; do
    python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/b.err | tail -1 > gpurun_out/ab_${v}_$w.json
    python -c "
import json; d=json.load(open('gpurun_out/ab_${v}_$w.json')); print('$v $w', d['ms_per_step'], {k:v['ms'] for k,v in d['stages'].items()})"
  done
done
unset WGS_RASTERIZER_LIB
tail -5 gpurun_out/b.err
