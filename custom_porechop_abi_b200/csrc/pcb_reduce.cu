/*
 * pcb_reduce.cu -- device-side reductions of the cross-product result stream.
 *
 * The DP kernels leave one 32-byte record per (window, adapter) pair in HBM.  The reference's callers consume far less:
 *   phase A (porechop_abi.py:200-245 -> nanopore_read.py:158-173): per adapter sequence the MAXIMUM full-adapter identity
 *           over all check reads;
 *   phase B (porechop_abi.py:359-438 -> nanopore_read.py:175-217): only the pairs whose aligned-region identity is above
 *           end_threshold (and that pass the edge / min_trim_size tests), plus -- for barcode sequences -- the full-adapter
 *           identity of every pair (nanopore_read.py:190-195, :212-217).
 * These kernels turn the records into exactly that, so that a few percent of the bytes cross PCIe.  Integer arithmetic
 * only: an identity is compared as the rational it is.
 */
#include "pcb_kernels.h"

#include <algorithm>
#include <cuda_runtime.h>
#include <stdint.h>

namespace {

// ---- phase A: best identity per adapter -------------------------------------------------------------------------------
// key of a record: floor(matches * 2^36 / full_len) in the high bits (two different ratios with lengths < 4096 differ by
// more than 2^-24, so the floor keeps their order), then matches and full_len themselves.  0 = no alignment.
__device__ __forceinline__ unsigned long long best_key(const pcb_result &r)
{
    if (r.read_start == -1 || r.full_len <= 0 || r.full_len >= 4096 || r.matches < 0 || r.matches >= 4096) return 0ull;
    const unsigned long long q = ((unsigned long long)r.matches << 36) / (unsigned long long)r.full_len;
    return (q << 24) | ((unsigned long long)r.matches << 12) | (unsigned long long)r.full_len;
}

__global__ void k_best(const pcb_result *__restrict__ rec, int64_t n_windows, int32_t n_ids, unsigned long long *__restrict__ keys)
{
    // block (x = window chunk, y = adapter): max over its windows, one atomic per block
    const int k = blockIdx.y;
    unsigned long long best = 0ull;
    for (int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; w < n_windows; w += (int64_t)gridDim.x * blockDim.x) {
        const unsigned long long key = best_key(rec[w * n_ids + k]);
        best = key > best ? key : best;
    }
    for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, d); best = o > best ? o : best; }
    __shared__ unsigned long long s[32];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x < 32) {
        best = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0ull;
        for (int d = 16; d > 0; d >>= 1) { const unsigned long long o = __shfl_xor_sync(0xffffffffu, best, d); best = o > best ? o : best; }
        if (threadIdx.x == 0 && best) atomicMax(keys + k, best);
    }
}

__global__ void k_best_decode(const unsigned long long *__restrict__ keys, int32_t n_ids, pcb_best *__restrict__ out)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_ids) return;
    const unsigned long long key = keys[k];
    pcb_best b; b.matches = (int32_t)((key >> 12) & 0xFFFull); b.full_len = (int32_t)(key & 0xFFFull);
    out[k] = b;
}

// ---- phase B: hits + barcode identities --------------------------------------------------------------------------------
// nanopore_read.py:187-188 / :209-210 with the identity the reference parses from "%f": float("%f" % x) > threshold
// <=> x >= (floor(threshold * 10^6) + 1/2) * 10^-6 (x = 100 * matches / aligned_len has no tie at six decimals for
// lengths < 512; '>=' keeps a tie on the safe side: the caller re-checks the few records it gets).
__device__ __forceinline__ bool hit_pass(const pcb_result &r, const pcb_hit_filter &f)
{
    const bool failed = r.read_start == -1;
    if (!failed && r.aligned_len <= 0) return false;                     // "-nan" compares false
    const long long num = failed ? 0ll : 200000000ll * (long long)r.matches, den = failed ? 1ll : (long long)r.aligned_len;
    if (num < (2ll * f.threshold_e6 + 1ll) * den) return false;
    const int rs = r.read_start, re = failed ? 0 : r.read_end + 1;      // read_end as align_adapter returns it (exclusive)
    if (f.at_end ? rs == 0 : re == f.end_size) return false;
    return re - rs >= f.min_trim_size;
}

// one warp per window: hits of the window
__global__ void k_hits_count(const pcb_result *__restrict__ rec, int64_t n_windows, int32_t n_ids, pcb_hit_filter f, int32_t *__restrict__ cnt)
{
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_windows) return;
    const int lane = threadIdx.x & 31;
    int n = 0;
    for (int k0 = 0; k0 < n_ids; k0 += 32) {
        const int k = k0 + lane;
        const bool p = k < n_ids && hit_pass(rec[w * n_ids + k], f);
        n += __popc(__ballot_sync(0xffffffffu, p));
    }
    if (lane == 0) cnt[w] = n;
}

// one warp per window: the trim amount its kept alignments ask for (find_start_trim / find_end_trim, nanopore_read.py:187-190,
// :209-212, before extra_trim_size is added): max of read_end (exclusive) at the start, of end_size - read_start at the end;
// -1 = no alignment kept
__global__ void k_trims(const pcb_result *__restrict__ rec, int64_t n_windows, int32_t n_ids, pcb_hit_filter f, int32_t *__restrict__ trim)
{
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_windows) return;
    const int lane = threadIdx.x & 31;
    int best = -1;
    for (int k = lane; k < n_ids; k += 32) {
        const pcb_result r = rec[w * n_ids + k];
        if (hit_pass(r, f)) best = max(best, f.at_end ? f.end_size - r.read_start : r.read_end + 1);
    }
    best = __reduce_max_sync(0xffffffffu, best);
    if (lane == 0) trim[w] = best;
}

// exclusive prefix sum of int32 counts, three steps (1024 per block)
__global__ void k_scan_blocks(int32_t *__restrict__ v, int64_t n, int32_t *__restrict__ bsum)
{
    __shared__ int s[32];
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int x = i < n ? v[i] : 0;
    int inc = x;
    for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
    if (lane == 31) s[wid] = inc;
    __syncthreads();
    if (wid == 0) {
        int t = s[lane];
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, t, d); if (lane >= d) t += o; }
        s[lane] = t;
    }
    __syncthreads();
    const int before = wid ? s[wid - 1] : 0;
    if (i < n) v[i] = before + inc - x;
    if (threadIdx.x == 1023) bsum[blockIdx.x] = before + inc;
}
__global__ void k_scan_sums(int32_t *__restrict__ bsum, int64_t nb, long long *__restrict__ total)
{
    // one block; serial over tiles of 1024 block sums
    __shared__ int s[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int64_t t0 = 0; t0 < nb; t0 += 1024) {
        const int64_t i = t0 + threadIdx.x;
        const int x = i < nb ? bsum[i] : 0;
        int inc = x;
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += o; }
        if (lane == 31) s[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int t = s[lane];
            for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, t, d); if (lane >= d) t += o; }
            s[lane] = t;
        }
        __syncthreads();
        const int before = wid ? s[wid - 1] : 0;
        if (i < nb) bsum[i] = (int)(carry + before + inc - x);
        __syncthreads();
        if (threadIdx.x == 1023) carry += before + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}
__global__ void k_scan_add(int32_t *__restrict__ v, int64_t n, const int32_t *__restrict__ bsum)
{
    const int64_t i = (int64_t)blockIdx.x * 1024 + threadIdx.x;
    if (i < n) v[i] += bsum[blockIdx.x];
}

// one warp per window: writes the window's hits at its slot range, in adapter order (the order in which the reference
// appends them), and the identity words of the wanted adapters
__global__ void k_hits_fill(const pcb_result *__restrict__ rec, int64_t n_windows, int32_t n_ids, pcb_hit_filter f,
                            const int32_t *__restrict__ first, pcb_hit *__restrict__ hits, int64_t cap,
                            const int32_t *__restrict__ ident_col, int32_t n_want, uint16_t *__restrict__ ident, int32_t *err_flag)
{
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n_windows) return;
    const int lane = threadIdx.x & 31;
    int64_t at = first[w];
    for (int k0 = 0; k0 < n_ids; k0 += 32) {
        const int k = k0 + lane;
        pcb_result r;
        bool p = false;
        if (k < n_ids) { r = rec[w * n_ids + k]; p = hit_pass(r, f); }
        const uint32_t m = __ballot_sync(0xffffffffu, p);
        if (p) {
            const int64_t slot = at + __popc(m & ((1u << lane) - 1u));
            if (slot < cap) {
                pcb_hit h;
                h.window = (int32_t)w; h.id_pos = k; h.read_start = r.read_start; h.read_end = r.read_end; h.raw_score = r.raw_score;
                h.matches = r.matches; h.aligned_len = r.aligned_len; h.full_len = r.full_len;
                hits[slot] = h;
            }
        }
        at += __popc(m);
        if (k < n_ids && ident && ident_col[k] >= 0) {
            uint16_t v = 0;
            if (r.read_start != -1) {
                if (r.matches >= 128 || r.full_len >= 512 || r.matches < 0 || r.full_len < 0) atomicOr(err_flag, 32);
                v = (uint16_t)((r.matches << 9) | r.full_len);
            }
            ident[w * n_want + ident_col[k]] = v;
        }
    }
}

// ---- barcode binning: the two best barcode identities of every window ---------------------------------------------------
// determine_barcode (nanopore_read.py:408-482) sorts a read's barcode scores in descending order (stable: the first in
// adapter order among equals) and looks at the first two.  One thread per window over its identity words.
__device__ __forceinline__ bool ident_gt(uint32_t a, uint32_t b)      // identity a > identity b, exactly (0 = failed = 0.0)
{
    const uint32_t ma = a >> 9, la = a & 0x1FFu, mb = b >> 9, lb = b & 0x1FFu;
    return ma * (lb ? lb : 1u) * (la ? 1u : 0u) > mb * (la ? la : 1u) * (lb ? 1u : 0u);
}
__global__ void k_top2(const uint16_t *__restrict__ ident, int64_t n_windows, int32_t n_want, pcb_top2 *__restrict__ out)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_windows) return;
    int b1 = -1, b2 = -1; uint32_t v1 = 0, v2 = 0;
    for (int j = 0; j < n_want; ++j) {
        const uint32_t v = ident[w * n_want + j];
        if (b1 < 0 || ident_gt(v, v1)) { b2 = b1; v2 = v1; b1 = j; v1 = v; }
        else if (b2 < 0 || ident_gt(v, v2)) { b2 = j; v2 = v; }
    }
    pcb_top2 t; t.best = (int16_t)b1; t.second = (int16_t)b2; t.best_ident = (uint16_t)v1; t.second_ident = (uint16_t)v2;
    out[w] = t;
}

}  // namespace

cudaError_t pcb_launch_trims(const pcb_result *rec, int64_t n_windows, int32_t n_ids, const pcb_hit_filter &f, int32_t *trim, cudaStream_t st)
{
    if (n_windows <= 0) return cudaSuccess;
    k_trims<<<(unsigned)((n_windows * 32 + 255) / 256), 256, 0, st>>>(rec, n_windows, n_ids, f, trim);
    return cudaGetLastError();
}

cudaError_t pcb_launch_top2(const uint16_t *ident, int64_t n_windows, int32_t n_want, pcb_top2 *out, cudaStream_t st)
{
    if (n_windows <= 0) return cudaSuccess;
    k_top2<<<(unsigned)((n_windows + 127) / 128), 128, 0, st>>>(ident, n_windows, n_want, out);
    return cudaGetLastError();
}

cudaError_t pcb_launch_best(const pcb_result *rec, int64_t n_windows, int32_t n_ids, unsigned long long *keys, pcb_best *out, cudaStream_t st)
{
    if (n_ids <= 0) return cudaSuccess;
    cudaError_t e = cudaMemsetAsync(keys, 0, sizeof(unsigned long long) * n_ids, st);
    if (e != cudaSuccess) return e;
    if (n_windows > 0) {
        const unsigned gx = (unsigned)std::max<int64_t>(1, std::min<int64_t>(64, (n_windows + 255) / 256));
        k_best<<<dim3(gx, (unsigned)n_ids), 256, 0, st>>>(rec, n_windows, n_ids, keys);
    }
    k_best_decode<<<(n_ids + 127) / 128, 128, 0, st>>>(keys, n_ids, out);
    return cudaGetLastError();
}

// cnt: n_windows + 1 ints (becomes the first hit slot of every window); bsum: ceil(n_windows / 1024) + 1 ints
cudaError_t pcb_launch_hits(const pcb_result *rec, int64_t n_windows, int32_t n_ids, const pcb_hit_filter &f, int32_t *cnt, int32_t *bsum,
                            long long *total, pcb_hit *hits, int64_t cap, const int32_t *ident_col, int32_t n_want, uint16_t *ident,
                            int32_t *err_flag, cudaStream_t st)
{
    if (n_windows <= 0 || n_ids <= 0) return cudaMemsetAsync(total, 0, sizeof(long long), st);
    const unsigned gw = (unsigned)((n_windows * 32 + 255) / 256);
    const int64_t nb = (n_windows + 1023) / 1024;
    k_hits_count<<<gw, 256, 0, st>>>(rec, n_windows, n_ids, f, cnt);
    k_scan_blocks<<<(unsigned)nb, 1024, 0, st>>>(cnt, n_windows, bsum);
    k_scan_sums<<<1, 1024, 0, st>>>(bsum, nb, total);
    k_scan_add<<<(unsigned)nb, 1024, 0, st>>>(cnt, n_windows, bsum);
    k_hits_fill<<<gw, 256, 0, st>>>(rec, n_windows, n_ids, f, cnt, hits, cap, ident_col, n_want, ident, err_flag);
    return cudaGetLastError();
}
