# usage (under gpurun): bash scripts/gpu_tests.sh [pytest args] -- the GPU suite
set -x
nvidia-smi --query-gpu=name,memory.total --format=csv
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 900 python -m pytest tests -m gpu -x -q "$@" 2>&1 | tee gpurun_out/pytest_last.log | tail -30
