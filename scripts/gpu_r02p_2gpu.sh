# round-2 run P (two GPUs): multi-GPU tests with the new default launch shapes (both paths)
set -x
timeout 600 python -m pytest tests/test_gpu_peer.py -m gpu -q -x 2>&1 | tail -4
