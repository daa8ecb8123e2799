# round-2 run U (one GPU): launch list of one window step of config 5's GROWN scene at 1 M rows (stage 30 of the schedule)
set -x
mkdir -p gpurun_out
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/r02u_c5_grown_1m_launches.csv python scripts/bench_config5.py --stop-at 31 --reps 1 --eager --streams 1 --profile-last > gpurun_out/ncu_u.log 2>&1
tail -2 gpurun_out/ncu_u.log | cut -c1-300
