# usage (under gpurun): bash scripts/gpu_e2e.sh <tag> -- GPU tests, then the ont bench line with the host-side time line of an e2e step
set -x
TAG=${1:-x}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_$TAG.log
PCB_E2E_TRACE=1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ont_$TAG.json 2> gpurun_out/bench_ont_$TAG.err
grep "e2e step" gpurun_out/bench_ont_$TAG.err; tail -3 gpurun_out/bench_ont_$TAG.err
python - <<PY
import json
d = json.loads(open('gpurun_out/bench_ont_$TAG.json').read().strip().split('\n')[-1])
print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'd2h', d['e2e']['d2h_bytes_per_step'], 'roofline', d['roofline']['frac'], 'phase_e2e', d.get('phase_e2e', {}).get('value'))
PY
