# usage (under gpurun): PCB_SO=... bash scripts/gpu_one.sh <tag> [n]  -- instruction count / issue rate of the 24-base trace kernel
TAG=${1:-x}; N=${2:-65536}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python scripts/exp/one.py $N > gpurun_out/one_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio,l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum,l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --kernel-name-base demangled -k 'regex:k_fast<\(bool\)1, \(int\)24, \(bool\)1' -s 1 -c 1 --csv --log-file gpurun_out/one_$TAG.csv python scripts/exp/one.py $N > gpurun_out/one_ncu_$TAG.log 2>&1
echo rc=$?; cat gpurun_out/one_$TAG.csv | tail -14
