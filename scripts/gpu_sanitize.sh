# usage (under gpurun): bash scripts/gpu_sanitize.sh <tag>  -- compute-sanitizer memcheck + racecheck on the trace, long-read and reduction tests
set -x
TAG=${1:-r2}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
SEL='trace_queue or long_reads_scan or reduced_outputs or windows_cross_product'
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "$SEL" > gpurun_out/sanitizer_${tool}_$TAG.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer_${tool}_$TAG.log | tail -4
done
