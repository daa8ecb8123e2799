# round-2 run I (two GPUs): bench.py at N=2, e2e with the upload allgather on the copy stream
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29551 bench.py --gpus 2 --steps 30 --warmup 5 --no-window > gpurun_out/bench_n2i.json 2> gpurun_out/bench_n2i.err; tail -5 gpurun_out/bench_n2i.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2i.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['h2d_bytes_per_step'])
PY
