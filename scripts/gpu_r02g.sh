# round-2 run G (one GPU): suite on the in-tree library (queued preprocess_bwd), strip stage times, render_bwd variant E
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -8
python scripts/strip_stages.py --steps 10 2>&1 | tail -12
for v in base bwde; do
  if [ $v = base ]; then unset WGS_RASTERIZER_LIB; else export WGS_RASTERIZER_LIB=$PWD/scripts/variants/libwgs_$v.so; fi
  if [ $v != base ]; then python -m pytest tests/test_gpu_api.py tests/test_gpu_parity_golden.py tests/test_gpu_render.py -m gpu -q 2>&1 | tail -3; fi
  for w in config2 config3; do
    python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/g.err | tail -1 > gpurun_out/g_${v}_$w.json
    python -c "
import json; d=json.load(open('gpurun_out/g_${v}_$w.json')); print('$v $w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
  done
done
unset WGS_RASTERIZER_LIB
python scripts/strip_stages.py --steps 5 --only 0,3,atomic | tail -1
WGS_RASTERIZER_LIB=$PWD/scripts/variants/libwgs_bwde.so python scripts/strip_stages.py --steps 5 --only 0,3,atomic | tail -1
tail -5 gpurun_out/g.err
