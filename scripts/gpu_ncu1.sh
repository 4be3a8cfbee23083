# usage (under gpurun): bash scripts/gpu_ncu1.sh <tag> list|trace|scan  -- ONE ncu pass per call, after the same command ran clean
set -x
TAG=${1:-r1}; WHAT=${2:-list}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --reads-per-gpu 32768 --no-cpu-baseline"
$CMD > gpurun_out/ncu_plain_$TAG.log 2>&1 || { echo "plain run failed"; exit 1; }
case $WHAT in
list)  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1 ;;
trace) ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_fast<\(bool\)1, \(int\)24, \(bool\)1' -s 2 -c 1 -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1 ;;
scan)  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_fast<\(bool\)0' -s 1 -c 1 -f -o gpurun_out/prof_${TAG}_scan $CMD > gpurun_out/ncu_full2_$TAG.log 2>&1 ;;
esac
echo "ncu $WHAT rc=$?"
ls -la gpurun_out | tail -4
