# usage (under gpurun): bash scripts/gpu_bench.sh <tag> [workloads...]  -- smoke, GPU tests, then the bench line of every workload
set -x
TAG=${1:-r2}; shift
WL=${@:-ont ultralong barcodes}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_$TAG.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
for w in $WL; do
  timeout 1200 python bench.py --workload $w > gpurun_out/bench_${w}_$TAG.json 2> gpurun_out/bench_${w}_$TAG.err; echo "bench $w rc=$?"
  cat gpurun_out/bench_${w}_$TAG.json; tail -5 gpurun_out/bench_${w}_$TAG.err
done
