# N=8 box: mapping windows (config 3: 10 keyframes / 1 M Gaussians; config 5's shape: 16 keyframes / 2 M Gaussians),
# whole rank step as one CUDA graph; the eager variant only if the captured one fails
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
run() { # name, args...
  n=$1; shift
  timeout 120 $TR --master-port 29522 scripts/bench_window.py --steps 10 --warmup 3 "$@" 2>>gpurun_out/w8.err | tail -1 > gpurun_out/$n.json
  python -c "import json,sys;d=json.load(open(sys.argv[1]));print(sys.argv[1],d['value'],d['ms_per_step'])" gpurun_out/$n.json
}
run w8_c5_peer_graph --workload config5 --peer auto --graph || run w8_c5_peer --workload config5 --peer auto
run w8_c3_strips_peer_graph --peer auto --strips --graph || run w8_c3_strips_peer --peer auto --strips
tail -3 gpurun_out/w8.err
