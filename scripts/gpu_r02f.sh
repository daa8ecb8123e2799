# round-2 run F (one GPU): ncu of preprocess_bwd / preprocess_fwd / scatter in strip mode; window with the fused loss
set -x
mkdir -p gpurun_out
python scripts/strip_stages.py --steps 3 --only 0,3,atomic | tail -1
ncu --set full --clock-control none --import-source on -k regex:"preprocess_|scatter" -s 12 -c 3 -f -o gpurun_out/r02f_strip python scripts/strip_stages.py --steps 3 --only 0,3,atomic > gpurun_out/ncu_strip.log 2>&1
tail -2 gpurun_out/ncu_strip.log
python scripts/bench_window.py --steps 10 --warmup 3 --parity 2>>gpurun_out/f.err | tail -1 | cut -c1-1200
python scripts/bench_window.py --steps 10 --warmup 3 --emulate 8 2>>gpurun_out/f.err | tail -1 | cut -c1-300
python bench.py --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_ours_f.json 2>>gpurun_out/f.err; cut -c1-1500 gpurun_out/bench_ours_f.json
tail -5 gpurun_out/f.err
