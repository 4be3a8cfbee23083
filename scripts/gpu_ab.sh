# usage (under gpurun): bash scripts/gpu_ab.sh <tag> -- parity tests, then main build against every variant in scratch_dbg/ (kernel timing + bench line)
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$TAG.log
cat > /tmp/line.py <<'PY'
import json, sys
d = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('bench ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'])
PY
( for so in main scratch_dbg/*.so; do
    [ "$so" = main ] && unset PCB_SO || { [ -f "$so" ] || continue; export PCB_SO=$PWD/$so; }
    echo "== $so"
    timeout 300 python scripts/exp/kbench.py 131072
    timeout 300 python scripts/exp/kbench_long.py 16384 rapid
    timeout 300 python scripts/exp/kbench_long.py 16384 native
    timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python /tmp/line.py
  done ) > gpurun_out/ab_$TAG.log 2>&1
cat gpurun_out/ab_$TAG.log
