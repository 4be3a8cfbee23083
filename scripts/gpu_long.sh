set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out; TAG=r2s
python scripts/exp/kbench_long.py 16384
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 3 -c 3 --csv --log-file gpurun_out/long_fast_$TAG.csv python scripts/exp/kbench_long.py 16384 > /dev/null 2>&1
PCB_NO_FAST=1 python scripts/exp/kbench_long.py 16384
PCB_NO_FAST=1 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast|k_rounds|k_align' -s 1 -c 1 --csv --log-file gpurun_out/long_general_$TAG.csv python scripts/exp/kbench_long.py 16384 > /dev/null 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ont_$TAG.json 2>gpurun_out/bench_ont_$TAG.err; tail -c 900 gpurun_out/bench_ont_$TAG.json
