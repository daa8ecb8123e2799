# round-2 run C (one GPU): whole GPU suite on the compositing-kernel variants, ncu --set full of both render kernels
set -x
mkdir -p gpurun_out
python -m pytest tests/test_gpu_arena.py -m gpu -x -q 2>&1 | tail -5
export WGS_RASTERIZER_LIB=$PWD/scripts/variants/libwgs_both.so
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-window > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"render_" -s 16 -c 2 -f -o gpurun_out/r02c_render python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-window > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
unset WGS_RASTERIZER_LIB
# where does the time of ONE rank of an 8-rank window step go (eager, one stream, rank 0)
python scripts/bench_window.py --emulate 8 --emulate-rank 0 --eager --streams 1 --steps 3 --warmup 2 | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02c_rank0_launches.csv python scripts/bench_window.py --emulate 8 --emulate-rank 0 --eager --streams 1 --steps 2 --warmup 1 > gpurun_out/ncu_rank0.log 2>&1
tail -2 gpurun_out/ncu_rank0.log | cut -c1-300
# the in-tree library (forward variant C + programmatic dependent launch): whole suite, then PDL on / off
python -m pytest tests -m gpu -x -q 2>&1 | tail -8
for pdl in 1 0; do for w in config2 config3; do
  WGS_PDL=$pdl python bench.py --workload $w --steps 30 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/c.err | tail -1 > gpurun_out/pdl${pdl}_$w.json
  python -c "
import json; d=json.load(open('gpurun_out/pdl${pdl}_$w.json')); print('pdl=$pdl $w', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
done; done
for pdl in 1 0; do WGS_PDL=$pdl python scripts/bench_window.py --steps 10 --warmup 3 2>>gpurun_out/c.err | tail -1 | cut -c1-200; done
tail -5 gpurun_out/c.err
