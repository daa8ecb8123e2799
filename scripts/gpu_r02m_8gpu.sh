# round-2 run M (eight GPUs): bench.py at N=8 again -- e2e upload with the one-block rendezvous, window plan picked by timing
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29571 bench.py --gpus 8 --steps 30 --warmup 5 --window-steps 20 > gpurun_out/bench_n8m.json 2> gpurun_out/bench_n8m.err; tail -2 gpurun_out/bench_n8m.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n8m.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['h2d_bytes_per_step'], d['window']['value'], d['window']['config']['plan'], d['window']['config']['plan_trials_ms'], d['multi_gpu_parity']['ok'], d['multi_gpu_parity']['max_rel_err'])
PY
timeout 300 $TR4 --master-port 29572 bench.py --gpus 4 --steps 30 --warmup 5 --window-steps 20 > gpurun_out/bench_n4m.json 2> gpurun_out/bench_n4m.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n4m.json').read().strip().splitlines()[-1])
print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value'], d['window']['value'], d['window']['config']['plan'], d['multi_gpu_parity']['ok'])
PY
