# usage (under gpurun): bash scripts/gpu_ncu.sh <tag> [kernel regex] [skip]  -- launch list + one full capture of the dominant kernel
set -x
TAG=${1:-r1}
KRE=${2:-k_fast}
SKIP=${3:-2}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --reads-per-gpu 8192 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
$CMD > gpurun_out/ncu_plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$KRE -s $SKIP -c 2 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
