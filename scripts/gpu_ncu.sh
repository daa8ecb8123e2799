# ncu --set full capture of kernels matching $KREGEX (after a plain run of the same command exits 0)
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-render_}" -s ${SKIP:-16} -c ${COUNT:-2} -f -o gpurun_out/prof python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
