# round-2 run R (one GPU): config 5 at its final size -- the whole step on one GPU and the per-rank compute of an 8-rank step
set -x
mkdir -p gpurun_out
python scripts/bench_config5.py --fixed-rows 2000000 --reps 5 2>gpurun_out/r.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('n1', d['all_steps_ms'])"
python scripts/bench_config5.py --fixed-rows 2000000 --reps 5 --emulate 8 2>>gpurun_out/r.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('rank0 of 8', d['all_steps_ms'])"
python scripts/bench_config5.py --fixed-rows 2000000 --reps 5 --emulate 8 --streams 1 2>>gpurun_out/r.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('rank0 of 8, 1 stream', d['all_steps_ms'])"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02r_c5_rank0_launches.csv python scripts/bench_config5.py --fixed-rows 2000000 --reps 1 --emulate 8 --streams 1 --eager > gpurun_out/ncu_r.log 2>&1
tail -2 gpurun_out/ncu_r.log | cut -c1-200
tail -3 gpurun_out/r.err
