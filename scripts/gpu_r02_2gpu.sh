# round-2, two GPUs: the multi-GPU tests (peer allreduce kernel, WindowTrainer with the real rasterizer), then bench.py at N=2
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer.py -m gpu -x -q -rs 2>&1 | tail -15 | tee gpurun_out/pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -3 gpurun_out/bench_n2.err; cut -c1-2500 gpurun_out/bench_n2.json
