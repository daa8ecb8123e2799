# One GPU-box round: parity tests, smoke, both bench arms, ncu launch list + full capture of the render kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -3
if [ -z "$SKIP_REF" ]; then
python bench.py --impl reference --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; tail -3 gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
fi
python bench.py --steps 100 --warmup 10 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; tail -3 gpurun_out/bench_ours.err; cat gpurun_out/bench_ours.json
python scripts/host_profile.py 2>&1 | head -14
if [ -n "$NCU" ]; then
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'preprocess|tile_|scatter|render_' -s 32 -c 8 -f -o gpurun_out/prof python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
fi
