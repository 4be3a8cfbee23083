# usage (under gpurun): bash scripts/gpu_gens2.sh <tag> -- two-stripe classes with 2..4 generations (PCB_GENS_TWO)
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
PCB_GENS_TWO=4 timeout 600 python -m pytest tests -m gpu -x -q -k "two_stripe or long_adapters or no_alloc" > gpurun_out/pytest_two4_$TAG.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_two4_$TAG.log
( for g in 2 3 4; do echo "== PCB_GENS_TWO=$g"; PCB_GENS_TWO=$g timeout 300 python scripts/exp/kbench_long.py 16384 rapid; done ) > gpurun_out/gens2_$TAG.log 2>&1
cat gpurun_out/gens2_$TAG.log
