cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
make -C oracle libporechop_oracle.so >/dev/null
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in PCB_STREAMS=4 PCB_STREAMS=6 PCB_STREAMS=8; do
  echo "== $v"
  env $v python scripts/exp/kbench.py 131072 2>/dev/null | head -2
  env $v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-phase-e2e > gpurun_out/bench_x.json; python -c "
import json; d=json.load(open('gpurun_out/bench_x.json')); print('ms/step %.2f e2e %.0f roof %.4f'%(d['ms_per_step'], d['e2e']['value'], d['roofline']['frac']))"
done
