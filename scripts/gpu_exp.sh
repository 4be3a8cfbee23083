# usage (under gpurun): bash scripts/gpu_exp.sh <tag> [n_reads] -- parity tests, then kernel timing for the main build and any variant in scratch_dbg/
set -x
TAG=${1:-x}
N=${2:-32768}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
make -C oracle libporechop_oracle.so > /dev/null
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_$TAG.log
( echo "== main"; timeout 300 python scripts/exp/kbench.py $N
  for mb in 1024 3072 12288; do echo "== main PCB_SCRATCH_MB=$mb"; PCB_SCRATCH_MB=$mb timeout 300 python scripts/exp/kbench.py $N 24; done
  for so in scratch_dbg/*.so; do [ -f "$so" ] && { echo "== $so"; PCB_SO=$PWD/$so timeout 300 python scripts/exp/kbench.py $N 24; }; done
  ) > gpurun_out/exp_$TAG.log 2>&1
cat gpurun_out/exp_$TAG.log
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/kl_$TAG.csv python scripts/exp/kbench.py $N 24 > gpurun_out/kl_$TAG.log 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/kl_$TAG.csv')) if len(r)>5]
hdr=rows[0]; ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ui=hdr.index('Metric Unit')
for r in rows[1:]:
    print(r[ki][:60], r[vi], r[ui])
PY
