# round-2 run T (one GPU): config 5 as stated at N=1 on the final code, launch list of one rank's step at 2 M, window shapes
set -x
mkdir -p gpurun_out
timeout 900 python scripts/bench_config5.py --reps 3 > gpurun_out/c5_n1t.json 2> gpurun_out/c5_n1t.err; tail -2 gpurun_out/c5_n1t.err; cut -c1-700 gpurun_out/c5_n1t.json
python scripts/bench_window.py --workload config5 --steps 10 --warmup 3 2>>gpurun_out/t.err | tail -1 > gpurun_out/win_c5shape_n1t.json; cut -c1-200 gpurun_out/win_c5shape_n1t.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02t_c5_rank0_launches.csv python scripts/bench_config5.py --fixed-rows 2000000 --reps 1 --emulate 8 --streams 1 --eager > gpurun_out/ncu_t.log 2>&1
tail -2 gpurun_out/ncu_t.log | cut -c1-200
python bench.py --steps 50 --warmup 10 --window-steps 10 > gpurun_out/bench_ours_t.json 2>>gpurun_out/t.err; cut -c1-300 gpurun_out/bench_ours_t.json
