# round-2 run D (two GPUs): multi-GPU tests incl. sharded Adam, config 5 (small) at N=1 / N=2 replicated / N=2 sharded
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_peer.py -m gpu -x -q -rs 2>&1 | tail -15 | tee gpurun_out/pytest_2gpu.log
timeout 600 python scripts/bench_config5.py --small --reps 2 > gpurun_out/c5s_n1.json 2> gpurun_out/c5s_n1.err; tail -3 gpurun_out/c5s_n1.err
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
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
timeout 600 $TR --master-port 29523 bench.py --gpus 2 --steps 20 --warmup 5 --window-steps 10 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; tail -3 gpurun_out/bench_n2.err; cut -c1-1500 gpurun_out/bench_n2.json
