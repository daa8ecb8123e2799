# round-2 run W (eight GPUs), final code: config 5 as stated and its fixed-size shape
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29601 scripts/bench_config5.py --reps 3 > gpurun_out/c5_n8w.json 2> gpurun_out/c5_n8w.err; tail -2 gpurun_out/c5_n8w.err | cut -c1-300; cut -c1-600 gpurun_out/c5_n8w.json
timeout 240 $TR --master-port 29602 scripts/bench_window.py --steps 20 --warmup 5 --workload config5 > gpurun_out/win_c5shape_n8w.json 2> gpurun_out/win_n8w.err; cut -c1-300 gpurun_out/win_c5shape_n8w.json
timeout 300 $TR --master-port 29603 bench.py --gpus 8 --steps 30 --warmup 5 --window-steps 20 --no-cpu-baseline > gpurun_out/bench_n8w.json 2> gpurun_out/bench_n8w.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n8w.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['window']['value'], d['window']['config']['plan'], d['multi_gpu_parity']['ok'], d['multi_gpu_parity']['max_rel_err'])
PY
