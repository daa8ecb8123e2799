# N=8 box: peer allreduce check/timing, config-2 weak scaling (allreduce variants), config-3 mapping window
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29520 scripts/test_peer_allreduce.py 2>gpurun_out/n8.err | tail -1 > gpurun_out/peer_n8.json
cat gpurun_out/peer_n8.json
for m in nccl p2p nvls; do
  WGS_ALLREDUCE=$m $TR --master-port 29521 bench.py --gpus 8 --steps 50 --warmup 5 2>>gpurun_out/n8.err | tail -1 > gpurun_out/n8_$m.json
  python -c "import json;d=json.load(open('gpurun_out/n8_$m.json'));print('$m',d['value'],d['ms_per_step'],d['e2e']['value'])"
done
for args in "" "--strips" "--peer auto" "--peer auto --strips" "--peer p2p --strips"; do
  $TR --master-port 29522 scripts/bench_window.py --steps 10 --warmup 3 $args 2>>gpurun_out/n8.err | tail -1 > "gpurun_out/win_n8$(echo $args | tr -d ' -').json"
  python -c "import json,sys;d=json.load(open(sys.argv[1]));print(sys.argv[1],d['value'],d['ms_per_step'])" "gpurun_out/win_n8$(echo $args | tr -d ' -').json"
done
tail -3 gpurun_out/n8.err
