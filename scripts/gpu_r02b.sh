# round-2 run B (one GPU): parity suite incl. arena tests, config 5 as stated at N=1
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
timeout 900 python scripts/bench_config5.py --reps 3 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; tail -5 gpurun_out/c5_n1.err; cut -c1-1800 gpurun_out/c5_n1.json
