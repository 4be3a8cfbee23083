# usage (under gpurun): bash scripts/gpu_bench.sh <tag>
set -x
TAG=${1:-r1}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build(); g.smoke()" > gpurun_out/smoke_$TAG.log 2>&1; echo "smoke rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_$TAG.log
timeout 600 python bench.py --steps 2 --warmup 1 --reads-per-gpu 16384 --no-cpu-baseline > gpurun_out/bench_small_$TAG.json 2> gpurun_out/bench_small_$TAG.err; echo "bench small rc=$?"
cat gpurun_out/bench_small_$TAG.json; tail -5 gpurun_out/bench_small_$TAG.err
timeout 1200 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
cat gpurun_out/bench_$TAG.json; tail -5 gpurun_out/bench_$TAG.err
