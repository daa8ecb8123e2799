# round-2 run J (one GPU): the full round (tests, smoke, both bench arms, neighbouring benches, ncu launch list + full
# capture) and config 5 as stated at N=1 on the final code
NCU=1 EXTRA=1 bash scripts/gpu_round.sh
set -x
timeout 900 python scripts/bench_config5.py --reps 3 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; tail -2 gpurun_out/c5_n1.err; cut -c1-700 gpurun_out/c5_n1.json
python scripts/bench_window.py --steps 10 --warmup 3 --emulate 8 2>>gpurun_out/j.err | tail -1 > gpurun_out/win_emulate8.json; cut -c1-300 gpurun_out/win_emulate8.json
