set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total --format=csv
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -20
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -30
