# round-2 run X (one GPU), final code: tests, smoke, bench (ours), ncu launch list + full capture
NCU=1 SKIP_REF=1 bash scripts/gpu_round.sh
