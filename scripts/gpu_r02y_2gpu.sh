# round-2 run Y (two GPUs), final code: the whole GPU suite with nothing skipped
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -rs 2>&1 | tail -6 | tee gpurun_out/pytest_2gpu.log
