# round-2, eight GPUs (one call, charged 8x): bench.py at N=8 (headline + config-3 window + multi_gpu_parity), config 5 as
# stated at N=8 (replicated and sharded Adam), window variants.  Every command under its own timeout.
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29531 bench.py --gpus 8 --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; tail -2 gpurun_out/bench_n8.err; cut -c1-600 gpurun_out/bench_n8.json
timeout 420 $TR --master-port 29532 scripts/bench_config5.py --reps 3 > gpurun_out/c5_n8.json 2> gpurun_out/c5_n8.err; tail -2 gpurun_out/c5_n8.err; cut -c1-900 gpurun_out/c5_n8.json
timeout 420 $TR --master-port 29533 scripts/bench_config5.py --reps 3 --sharded > gpurun_out/c5_n8_sharded.json 2> gpurun_out/c5_n8_sharded.err; tail -2 gpurun_out/c5_n8_sharded.err; cut -c1-900 gpurun_out/c5_n8_sharded.json
timeout 240 $TR --master-port 29534 scripts/bench_window.py --steps 20 --warmup 5 --parity > gpurun_out/win_n8_balanced.json 2> gpurun_out/win_n8.err; cut -c1-400 gpurun_out/win_n8_balanced.json
timeout 240 $TR --master-port 29535 scripts/bench_window.py --steps 20 --warmup 5 --plan strips > gpurun_out/win_n8_strips.json 2>> gpurun_out/win_n8.err; cut -c1-300 gpurun_out/win_n8_strips.json
timeout 240 $TR --master-port 29536 scripts/bench_window.py --steps 20 --warmup 5 --workload config5 > gpurun_out/win_c5shape_n8.json 2>> gpurun_out/win_n8.err; cut -c1-300 gpurun_out/win_c5shape_n8.json
tail -3 gpurun_out/win_n8.err
