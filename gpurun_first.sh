set -x
cd $GRAFT_REPO_ROOT
nvidia-smi --query-gpu=name,memory.total --format=csv
nproc
python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -40
