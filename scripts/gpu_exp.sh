# usage (under gpurun): bash scripts/gpu_exp.sh <tag>   -- parity tests, then kernel timing for the main build and any variant in scratch_dbg/
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_$TAG.log
( echo "== main"; timeout 300 python scripts/exp/kbench.py 16384
  for so in scratch_dbg/*.so; do [ -f "$so" ] && { echo "== $so"; PCB_SO=$PWD/$so timeout 300 python scripts/exp/kbench.py 16384; PCB_SO=$PWD/$so timeout 300 python scripts/exp/phases.py 32768 | head -3; }; done
  echo "== phases"; timeout 600 python scripts/exp/phases.py 32768 ) > gpurun_out/exp_$TAG.log 2>&1
cat gpurun_out/exp_$TAG.log
