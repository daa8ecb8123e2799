# round-2 run Q (eight GPUs), final code: bench.py at N=8, config 5 as stated (replicated / sharded Adam), config-5-shape window
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29591 bench.py --gpus 8 --steps 30 --warmup 5 --window-steps 20 > gpurun_out/bench_n8q.json 2> gpurun_out/bench_n8q.err; tail -2 gpurun_out/bench_n8q.err | cut -c1-300; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n8q.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['window']['value'], d['window']['config']['plan'], d['window']['config']['plan_trials_ms'], d['multi_gpu_parity']['ok'], d['multi_gpu_parity']['max_rel_err'])
PY
timeout 420 $TR --master-port 29592 scripts/bench_config5.py --reps 3 > gpurun_out/c5_n8q.json 2> gpurun_out/c5_n8q.err; tail -2 gpurun_out/c5_n8q.err | cut -c1-300; cut -c1-800 gpurun_out/c5_n8q.json
timeout 420 $TR --master-port 29593 scripts/bench_config5.py --reps 3 --sharded > gpurun_out/c5_n8q_sharded.json 2> gpurun_out/c5_n8q_sharded.err; tail -2 gpurun_out/c5_n8q_sharded.err | cut -c1-300; cut -c1-800 gpurun_out/c5_n8q_sharded.json
timeout 240 $TR --master-port 29594 scripts/bench_window.py --steps 20 --warmup 5 --workload config5 > gpurun_out/win_c5shape_n8q.json 2> gpurun_out/win_n8q.err; cut -c1-300 gpurun_out/win_c5shape_n8q.json
