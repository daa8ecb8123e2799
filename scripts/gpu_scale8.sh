# N=8 box: config-2 weak scaling bench + config-3 mapping window (whole views vs tile-row strips)
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$TR --master-port 29521 bench.py --gpus 8 --steps 50 --warmup 5 2>gpurun_out/n8.err | tail -1 > gpurun_out/n8.json
$TR --master-port 29522 scripts/bench_window.py --steps 10 --warmup 3 2>>gpurun_out/n8.err | tail -1 > gpurun_out/win_n8.json
$TR --master-port 29523 scripts/bench_window.py --steps 10 --warmup 3 --strips 2>>gpurun_out/n8.err | tail -1 > gpurun_out/win_n8_strips.json
tail -3 gpurun_out/n8.err
for f in n8 win_n8 win_n8_strips; do cut -c1-330 gpurun_out/$f.json; done
