# usage (under gpurun): bash scripts/gpu_checked.sh <tag>  -- the GPU tests against a -DPCF_CHECKED build (every scratch / shared-memory / result
# index of the trace kernels range-checked on the device; compute-sanitizer is closed on this pool)
set -x
TAG=${1:-r2}
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out scratch_dbg
make -C oracle libporechop_oracle.so > /dev/null
( cd custom_porechop_abi_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared -DPCF_CHECKED -o ../../scratch_dbg/checked.so pcb_kernels.cu pcb_fast.cu pcb_reduce.cu pcb_capi.cu ) 2>&1 | grep -E " error" 
PCB_SO=$PWD/scratch_dbg/checked.so timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/checked_$TAG.log 2>&1; echo "checked rc=$?"; tail -3 gpurun_out/checked_$TAG.log
