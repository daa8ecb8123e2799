# round-2 run Z (one GPU): ncu --set full of the sort kernels on config 5's grown scene at 1 M rows
set -x
mkdir -p gpurun_out
ncu --profile-from-start off --set full --clock-control none --import-source on -k regex:"tile_sort|scatter" -c 4 -f -o gpurun_out/r02z_sort python scripts/bench_config5.py --stop-at 31 --reps 1 --eager --streams 1 --profile-last > gpurun_out/ncu_z.log 2>&1
tail -2 gpurun_out/ncu_z.log | cut -c1-200
