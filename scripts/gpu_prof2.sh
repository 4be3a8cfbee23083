# usage (under gpurun): bash scripts/gpu_prof2.sh <tag> [n]  -- full ncu captures of pass 1 and of the four rounds launches (one per generation) (24-base exact class) on the kernel bench
set -x
TAG=${1:-x}; N=${2:-32768}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
CMD="python scripts/exp/kbench.py $N 24"
$CMD > gpurun_out/kb_$TAG.log 2>&1 || { echo "plain run failed"; cat gpurun_out/kb_$TAG.log; exit 1; }
cat gpurun_out/kb_$TAG.log
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_fast<\(int\)1, \(int\)24, \(bool\)1' -s 1 -c 1 -f -o gpurun_out/prof_${TAG}_p1 $CMD > gpurun_out/ncu_p1_$TAG.log 2>&1
echo "ncu p1 rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_rounds<\(int\)24, \(bool\)1' -s 4 -c 4 -f -o gpurun_out/prof_${TAG}_p2 $CMD > gpurun_out/ncu_p2_$TAG.log 2>&1
echo "ncu p2 rc=$?"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k 'regex:k_fast<\(int\)0' -s 1 -c 1 -f -o gpurun_out/prof_${TAG}_scan python bench.py --steps 1 --warmup 1 --reads-per-gpu 16384 --no-cpu-baseline --no-phase-e2e > gpurun_out/ncu_scan_$TAG.log 2>&1
echo "ncu scan rc=$?"
# the 72-row class against the general kernel on the same input: time, then instruction counts
python scripts/exp/kbench_long.py 16384
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 3 -c 3 --csv --log-file gpurun_out/long_fast_$TAG.csv python scripts/exp/kbench_long.py 16384 > /dev/null 2>&1
PCB_NO_FAST=1 python scripts/exp/kbench_long.py 16384
PCB_NO_FAST=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 1 -c 1 --csv --log-file gpurun_out/long_general_$TAG.csv python scripts/exp/kbench_long.py 16384 > /dev/null 2>&1
ls -la gpurun_out | tail -5
