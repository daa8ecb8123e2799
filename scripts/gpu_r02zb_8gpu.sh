# round-2 run ZB (eight GPUs): config 5 as stated on the final sort kernels
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29611 scripts/bench_config5.py --reps 3 > gpurun_out/c5_n8zb.json 2> gpurun_out/c5_n8zb.err; tail -2 gpurun_out/c5_n8zb.err | cut -c1-300; cut -c1-400 gpurun_out/c5_n8zb.json
