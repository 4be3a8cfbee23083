# usage (under gpurun): bash scripts/gpu_static.sh <tag> -- the no-allocation entry point's test alone in a fresh process, then after the whole GPU suite
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_static_alone_$TAG.log 2>&1; echo "alone rc=$?"; tail -30 gpurun_out/pytest_static_alone_$TAG.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "suite rc=$?"; tail -8 gpurun_out/pytest_$TAG.log
