# round-2 run O (eight GPUs): launch-shape sweep of the peer exchange kernel
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
SWEEP_BLOCKS=${SWEEP_BLOCKS:-24,36,48,72,96} SWEEP_UNROLL=${SWEEP_UNROLL:-1,2} timeout 400 $TR --master-port 29581 scripts/tune_peer.py > gpurun_out/tune_peer_n8.json 2> gpurun_out/tune_peer_n8.err; tail -3 gpurun_out/tune_peer_n8.err | cut -c1-300; cat gpurun_out/tune_peer_n8.json
