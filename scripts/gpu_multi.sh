# usage (under gpurun --gpus N): bash scripts/gpu_multi.sh <tag> <N> [workloads...]  -- the sharded NCCL test, then the bench of every workload on N GPUs
set -x
TAG=${1:-r2}; N=${2:-2}; shift; shift
WL=${@:-ont ultralong barcodes}
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q -s > gpurun_out/sharded_N${N}_$TAG.log 2>&1; echo "sharded rc=$?"; tail -4 gpurun_out/sharded_N${N}_$TAG.log
for w in $WL; do
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29555 bench.py --gpus $N --workload $w --steps 5 --warmup 3 > gpurun_out/bench_${w}_N${N}_$TAG.json 2> gpurun_out/bench_${w}_N${N}_$TAG.err; echo "bench $w N=$N rc=$?"
  cat gpurun_out/bench_${w}_N${N}_$TAG.json | cut -c1-700; tail -3 gpurun_out/bench_${w}_N${N}_$TAG.err
done
