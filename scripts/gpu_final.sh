# usage (under gpurun): bash scripts/gpu_final.sh <tag>  -- the evidence of a round: tests, the default bench line, the launch list, full captures
set -x
TAG=${1:-r2}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
bash scripts/gpu_bench.sh $TAG ont
CMD="python bench.py --steps 1 --warmup 1 --reads-per-gpu 32768 --no-cpu-baseline --no-phase-e2e"
$CMD > gpurun_out/ncu_plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "ncu list rc=$?"
bash scripts/gpu_prof2.sh $TAG
