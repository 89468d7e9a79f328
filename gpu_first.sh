set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
nproc
python __graft_entry__.py smoke 2>&1 | tail -5
timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30
