# usage (under gpurun): bash scripts/gpu_ncu_two.sh <tag> -- one full ncu capture of generation 1 of the two-stripe rounds (rapid sets)
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
CMD="python scripts/exp/kbench_long.py 16384 rapid"
$CMD || exit 1
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rounds2' -s 3 -c 1 -f -o gpurun_out/prof_two_$TAG $CMD > gpurun_out/ncu_two_$TAG.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out | tail -3
