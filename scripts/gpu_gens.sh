# usage (under gpurun): bash scripts/gpu_gens.sh <tag> -- GPU tests, then pass 2 with 2..5 generations for the 24-base and 72-row classes (PCB_GENS)
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
for g in 5 4; do
  PCB_GENS=$g PCB_GENS_SMALL=$g timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_g${g}_$TAG.log 2>&1; echo "pytest gens=$g rc=$?"; tail -2 gpurun_out/pytest_g${g}_$TAG.log
done
cat > /tmp/line.py <<'PY'
import json, sys
d = json.loads(sys.stdin.read().strip().split('\n')[-1])
print('bench ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], 'checksum', d['e2e']['checksum'])
PY
( for g in 3 4 5; do
    echo "== PCB_GENS=$g"
    PCB_GENS=$g timeout 300 python scripts/exp/kbench.py 131072 24
    PCB_GENS=$g timeout 300 python scripts/exp/kbench_long.py 16384 native
    PCB_GENS=$g timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-phase-e2e 2>/dev/null | python /tmp/line.py
  done ) > gpurun_out/gens_$TAG.log 2>&1
cat gpurun_out/gens_$TAG.log
