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
if [ -n "$EXTRA" ]; then
# config 4 (3 M Gaussians, 1080p, SH 3) both arms; mapping windows at N=1 as one CUDA graph; prune_points
python bench.py --workload config4 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/c4_ours.json 2>> gpurun_out/extra.err; cut -c1-400 gpurun_out/c4_ours.json
python bench.py --impl reference --workload config4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/c4_ref.json 2>> gpurun_out/extra.err; cut -c1-400 gpurun_out/c4_ref.json
python scripts/bench_window.py --steps 10 --warmup 3 2>> gpurun_out/extra.err | tail -1 > gpurun_out/win_c3_n1_graph.json; cut -c1-200 gpurun_out/win_c3_n1_graph.json
python scripts/bench_window.py --workload config5 --steps 5 --warmup 2 2>> gpurun_out/extra.err | tail -1 > gpurun_out/win_c5_n1_graph.json; cut -c1-200 gpurun_out/win_c5_n1_graph.json
python scripts/bench_compact.py 2>> gpurun_out/extra.err | tail -1 > gpurun_out/compact.json; cat gpurun_out/compact.json
python scripts/bench_iteration.py 2>> gpurun_out/extra.err | tail -1 > gpurun_out/iteration.json; cut -c1-400 gpurun_out/iteration.json
python scripts/bench_iteration.py mapping 2>> gpurun_out/extra.err | tail -1 > gpurun_out/mapping_iteration.json; cut -c1-400 gpurun_out/mapping_iteration.json
python scripts/bench_iteration.py mapping_uncertainty 2>> gpurun_out/extra.err | tail -1 > gpurun_out/mapping_uncertainty_iteration.json; cut -c1-500 gpurun_out/mapping_uncertainty_iteration.json
python scripts/bench_ssim.py 2>> gpurun_out/extra.err | tail -1 > gpurun_out/ssim.json; cat gpurun_out/ssim.json
fi
if [ -n "$NCU" ]; then
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-window > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-window > gpurun_out/ncu.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'preprocess|tile_|scatter|render_' -s 32 -c 8 -f -o gpurun_out/prof python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-window > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
fi
