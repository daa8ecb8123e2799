# round-2 run E (one GPU): whole suite on the in-tree library (no -x), strip stage times
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q 2>&1 | tail -12
python scripts/strip_stages.py --steps 10 2>&1 | tail -14
