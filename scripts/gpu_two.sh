# usage (under gpurun): bash scripts/gpu_two.sh <tag> -- parity tests, then the two-stripe class (73..127 bases) against the general kernel
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 600 python -m pytest tests -m gpu -x -q -k "two_stripe or long_adapters" > gpurun_out/pytest_two_$TAG.log 2>&1; echo "pytest two rc=$?"; tail -25 gpurun_out/pytest_two_$TAG.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_$TAG.log
( timeout 300 python scripts/exp/kbench_long.py 16384 rapid
  PCB_NO_TWO=1 timeout 300 python scripts/exp/kbench_long.py 16384 rapid
  timeout 300 python scripts/exp/kbench_long.py 16384 native ) > gpurun_out/two_$TAG.log 2>&1
cat gpurun_out/two_$TAG.log
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 8 -c 4 --csv --log-file gpurun_out/two_fast_$TAG.csv python scripts/exp/kbench_long.py 16384 rapid > /dev/null 2>&1
PCB_NO_TWO=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 1 -c 1 --csv --log-file gpurun_out/two_general_$TAG.csv python scripts/exp/kbench_long.py 16384 rapid > /dev/null 2>&1
grep -h "k_fast\|k_rounds\|k_align" gpurun_out/two_fast_$TAG.csv gpurun_out/two_general_$TAG.csv | cut -c1-260
