# round-2 run H (two GPUs): parity suite after the scatter change, multi-GPU tests, bench.py at N=2 with the shared e2e
# upload, config 5 (small) at N=1 / N=2 replicated / N=2 sharded
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x -rs 2>&1 | tail -6 | tee gpurun_out/pytest_2gpu.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -5 gpurun_out/bench_n2.err; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['h2d_bytes_per_step'], d['window']['value'], d['multi_gpu_parity']['ok'], d['multi_gpu_parity']['max_rel_err'])
PY
timeout 600 python scripts/bench_config5.py --small --reps 2 > gpurun_out/c5s_n1.json 2> gpurun_out/c5s_n1.err; tail -3 gpurun_out/c5s_n1.err
timeout 600 $TR --master-port 29521 scripts/bench_config5.py --small --reps 2 > gpurun_out/c5s_n2.json 2> gpurun_out/c5s_n2.err; tail -3 gpurun_out/c5s_n2.err
timeout 600 $TR --master-port 29522 scripts/bench_config5.py --small --reps 2 --sharded > gpurun_out/c5s_n2_sharded.json 2> gpurun_out/c5s_n2_sharded.err; tail -3 gpurun_out/c5s_n2_sharded.err
python - <<'PY'
import json
for f in ("c5s_n1", "c5s_n2", "c5s_n2_sharded"):
    try:
        d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], d["value_at_final_size"], d["all_steps_ms"], [round(x, 4) for x in d["loss_sums"]])
    except Exception as e:
        print(f, "failed", e)
PY
python bench.py --workload config3 --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-window 2>>gpurun_out/h.err | tail -1 > gpurun_out/h_config3.json
python -c "
import json; d=json.load(open('gpurun_out/h_config3.json')); print('config3', d['ms_per_step'], d['value'], {k:v['ms'] for k,v in d['stages'].items()})"
