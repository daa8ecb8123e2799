# round-2 run L (two GPUs): peer tests (rendezvous), bench.py at N=2 with plan=auto and the rendezvous in the e2e upload
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_peer.py -m gpu -q -x 2>&1 | tail -4
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29561 bench.py --gpus 2 --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_n2l.json 2> gpurun_out/bench_n2l.err; tail -5 gpurun_out/bench_n2l.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2l.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['h2d_bytes_per_step'], d['window']['value'], d['window']['config']['plan'], d['window']['config']['plan_trials_ms'], d['multi_gpu_parity']['ok'], d['multi_gpu_parity']['max_rel_err'])
PY
