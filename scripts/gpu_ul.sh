cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for pp in 4 8 16; do
PCB_PLAN_PIECES=$pp PCB_E2E_TRACE=1 python bench.py --workload ultralong --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/ul_$pp.json 2> gpurun_out/ul_$pp.err
grep "e2e step" gpurun_out/ul_$pp.err | tail -1
python - <<PY
import json
d = json.loads(open('gpurun_out/ul_$pp.json').read().strip().split('\n')[-1])
print('pieces', $pp, 'ms_per_step', d['ms_per_step'], 'value', d['value'], 'e2e', d['e2e']['value'], 'checksum', d['e2e'].get('checksum'))
PY
done
