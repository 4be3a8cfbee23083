/*
 * pcb_kernels.cu -- sm_100a kernels of the adapter-alignment hot path.
 *
 * What is computed (reference: porechop_abi/src/adapter_align.cpp:11-31 = SeqAn 2.4.0 globalAlignment,
 * AlignConfig<true,true,true,true>, Score<int,Simple>, SingleTrace/GapsLeft; then
 * porechop_abi/src/alignment.cpp:5-110) is restated in DESIGN.md section 3; the tie-breaking rules the
 * flags below encode are those of seqan/align/dp_formula_affine.h:65-124 and dp_formula.h:147-163.
 *
 * How (B200-first, not a translation of SeqAn's column loop):
 *   - inter-task SIMD: ONE THREAD = TWO TASKS.  The two tasks of a thread live in the two signed
 *     16-bit halves of every 32-bit register and are advanced together with the native sm_100a
 *     packed-integer instructions VIADD.16x2, VIMNMX.S16x2, VIADDMNMX.S16x2 (DPX).  A warp = 64
 *     tasks that share one adapter, so the row loop, the adapter base (one broadcast LDS) and the
 *     control flow are warp-uniform.  No shuffles, no idle wavefront lanes.
 *   - the DP column (H of the left neighbour and M+gapOpen of the left neighbour, one register each
 *     per adapter row) lives in registers: 32 rows per stripe; longer adapters run as several stripes
 *     with the stripe boundary row spilled to a per-thread scratch line.
 *   - adapters are bottom-aligned in the 32-row register tile and the row loop is entered through a
 *     jump table, so the last adapter row is always register 31 and no per-row bounds test exists.
 *   - trace: 4 flag bits per cell, extracted from the sign bits of packed differences with two PRMT
 *     (sign-replicate mode) and bit-select LOP3s; 4 rows x 2 tasks x 4 bits = one 32-bit word, stored
 *     coalesced ([column][word][thread]) to an HBM scratch ring that only the owning thread reads back.
 *   - the scan variant (TRACE=false) is the same recurrence without flags: it finds the best end cell of
 *     long reads chunk by chunk; the trace variant is then run on a window around that cell.
 */
#include "pcb_kernels.h"
#include "align_stats.h"

#include <cuda_runtime.h>
#include <stdint.h>

namespace {

// ------------------------------------------------------------------------------------------------
// packed signed 16x2 helpers (all single native instructions on sm_100a, see profiles/sass_k_fast_r2.md)
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t vadd(uint32_t a, uint32_t b) { return __vadd2(a, b); }          // VIADD.16x2
__device__ __forceinline__ uint32_t vmax(uint32_t a, uint32_t b) { return __vmaxs2(a, b); }         // VIMNMX.S16x2
__device__ __forceinline__ uint32_t vaddmax(uint32_t a, uint32_t b, uint32_t c)                      // VIADDMNMX.S16x2
{
    return __viaddmax_s16x2(a, b, c);   // max(a + b, c) per half
}
// bytes of the result: sign(a.lo), sign(a.hi), sign(b.lo), sign(b.hi), each replicated to 0x00 / 0xFF
__device__ __forceinline__ uint32_t sign4(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, 0xfdb9;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t pk2(int v) { return ((uint32_t)v & 0xFFFFu) * 0x10001u; }
__device__ __forceinline__ int lo16(uint32_t x) { return (int)(short)(x & 0xFFFFu); }
__device__ __forceinline__ int hi16(uint32_t x) { return (int)x >> 16; }
__device__ __forceinline__ int half16(uint32_t x, int h) { return h ? hi16(x) : lo16(x); }

// SeqAn's char -> Dna5 table (seqan/basic/alphabet_residue_tabs.h:113-140); stored as code * PCB_CODE_SCALE
__device__ __forceinline__ uint8_t dna5_code(unsigned char c)
{
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1 * PCB_CODE_SCALE;
    case 'G': case 'g': return 2 * PCB_CODE_SCALE;
    case 'T': case 't': case 'U': case 'u': return 3 * PCB_CODE_SCALE;
    default: return 4 * PCB_CODE_SCALE;
    }
}

__global__ void k_encode(const char *__restrict__ ascii, uint8_t *__restrict__ codes, uint64_t n)
{
    // 16 bytes per thread, coalesced; tail handled bytewise
    const uint64_t i = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 16;
    if (i + 16 <= n && ((((uintptr_t)ascii) | ((uintptr_t)codes)) & 15) == 0) {
        uint4 v = *reinterpret_cast<const uint4 *>(ascii + i);
        uint32_t w[4] = { v.x, v.y, v.z, v.w };
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            uint32_t o = 0;
#pragma unroll
            for (int b = 0; b < 4; ++b) o |= (uint32_t)dna5_code((unsigned char)(w[k] >> (8 * b))) << (8 * b);
            w[k] = o;
        }
        *reinterpret_cast<uint4 *>(codes + i) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
        for (uint64_t k = i; k < n && k < i + 16; ++k) codes[k] = dna5_code((unsigned char)ascii[k]);
    }
}

// End windows cut out of the read blob on the device (nanopore_read.py:164,170,181,203: seq[:end_size] / seq[-end_size:]):
// the reads are in the task arena already (pcb_upload_reads), so neither a host-side gather nor a second copy is needed.
// One thread per (read, 4 bytes of its window record).
__global__ void k_gather_windows(const uint8_t *__restrict__ arena_t, const int64_t *__restrict__ roff, const int32_t *__restrict__ rlen,
                                 int64_t n, int end_size, int at_end, int stride, uint8_t *__restrict__ arena, int32_t *__restrict__ win_len)
{
    const int groups = stride / 4;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t w = idx / groups;
    if (w >= n) return;
    const int g = (int)(idx - w * groups);
    const int len = rlen[w], wl = len < end_size ? len : end_size;
    const uint8_t *src = arena_t + roff[w] + (at_end ? len - wl : 0);
    uint32_t word = 0;
#pragma unroll
    for (int b = 0; b < 4; ++b) { const int i = 4 * g + b; if (i < wl) word |= (uint32_t)src[i] << (8 * b); }
    reinterpret_cast<uint32_t *>(arena + w * stride)[g] = word;
    if (g == 0) win_len[w] = wl;
}

// order key of a tracked cell: larger score wins, then the cell SeqAn's scout visits first
// (dp_scout.h:167-180: strict '>' in visiting order).  ord = column for last-row cells, len + row for
// cells of the final column (dp_meta_info.h:205-216).  Two low bits carry V==M and H==M of that cell
// for _correctTraceValue (dp_algorithm_impl.h:1169-1185).
__device__ __forceinline__ int make_key(int score, int ord, int eqV, int eqH)
{
    return (int)(((uint32_t)score << 16) | ((uint32_t)(0x3FFF - ord) << 2) | (uint32_t)(eqV << 1) | (uint32_t)eqH);
}

}  // namespace

namespace {

struct HalfTask {            // one of the two tasks of a thread
    uint64_t off;
    int32_t  len, col0, m_true, forced, own0;
    int64_t  out_idx;        // < 0: no task in this half
};

__device__ __forceinline__ void fetch_task(const PcbSource &src, int64_t unit, int slot, int32_t &adapter, HalfTask &t)
{
    t.off = 0; t.len = 0; t.col0 = 0; t.m_true = 0; t.forced = -1; t.own0 = 0; t.out_idx = -1;
    if (src.units) {
        const PcbUnit u = src.units[unit];
        adapter = u.adapter;
        if (slot < u.ntask) {
            const PcbTask k = src.tasks[u.task0 + slot];
            t.off = k.off; t.len = k.len; t.col0 = k.col0; t.m_true = k.m_true; t.forced = k.forced_ord;
            t.own0 = k.own0; t.out_idx = k.out_idx;
        }
    } else {
        const int64_t g = unit / src.n_ids;
        const int32_t k = (int32_t)(unit - g * src.n_ids);
        adapter = src.ids[k];
        const int64_t w = g * 64 + slot;
        if (w < src.n_windows) {
            t.off = (uint64_t)w * (uint64_t)src.stride;
            t.len = src.win_len[w]; t.m_true = t.len;
            t.out_idx = w * src.n_ids_total + (src.id_pos ? src.id_pos[k] : k);
        }
    }
}

__device__ __forceinline__ void write_failed(pcb_result *out, int64_t idx)
{
    // the reference's "alignment failed" record (alignment.cpp:7-21; values as its build prints them)
    pcb_result r;
    r.read_start = -1; r.read_end = 0; r.adapter_start = -1; r.adapter_end = 0;
    r.raw_score = (int32_t)0x80000000; r.matches = 0; r.aligned_len = 0; r.full_len = 0;
    out[idx] = r;
}

// ------------------------------------------------------------------------------------------------
// One DP cell for the two tasks of this thread.  Register row R (compile-time).  Uses / updates:
//   Mo[R], nMo[R] (M + gapOpen of the left neighbour and its complement), Hl[R] (H of the left neighbour),
//   carries Mo_d (diagonal), Mo_u / nMo_u / Vu (cell above).
// Recurrence and ties: dp_formula_affine.h:65-124 (left operand of every _maxScore wins ties).
// ------------------------------------------------------------------------------------------------
#define PCB_CELL(R)                                                                                   \
    {                                                                                                 \
        const uint32_t ap = s_adp[R];                                                                 \
        const uint32_t t = __vminu2(rp ^ ap, 0x00010001u);          /* 1 = bases differ */            \
        const uint32_t Sp = ((t * 0xFFFFu) & XORC) ^ MATCHP;        /* score - gapOpen  */            \
        const uint32_t D = vadd(Mo_d, Sp);                                                            \
        const uint32_t Hn = vaddmax(Hl[R], GE, Mo[R]);                                                \
        const uint32_t Vn = vaddmax(Vu, GE, Mo_u);                                                    \
        const uint32_t G = vmax(Vn, Hn);                                                              \
        const uint32_t Mn = vmax(D, G);                                                               \
        const uint32_t MoN = vadd(Mn, GO);                                                            \
        if (TRACE) {                                                                                  \
            const uint32_t dh = vadd(vadd(Hl[R], GE1), ~Mo[R]);     /* <0: H opened   */              \
            const uint32_t dv = vadd(vadd(Vu, GE1), nMo_u);         /* <0: V opened   */              \
            const uint32_t dg = vadd(Hn, ~Vn);                      /* <0: V >= H     */              \
            const uint32_t dd = vadd(G, ~D);                        /* <0: diag >= gap */             \
            const uint32_t W = (sign4(dh, dv) & 0x55555555u) | (sign4(dg, dd) & 0xAAAAAAAAu);         \
            if (((R) & 3) == 0) Wacc = W;                                                             \
            else if (((R) & 3) == 1) Wacc = (Wacc & 0x03030303u) | (W & 0xFCFCFCFCu);                 \
            else if (((R) & 3) == 2) Wacc = (Wacc & 0x0F0F0F0Fu) | (W & 0xF0F0F0F0u);                 \
            else Wacc = (Wacc & 0x3F3F3F3Fu) | (W & 0xC0C0C0C0u);                                     \
            nMo_u = ~MoN;                                                                             \
        }                                                                                             \
        Mo_d = Mo[R]; Mo[R] = MoN; Hl[R] = Hn; Mo_u = MoN; Vu = Vn;                                   \
    }

#define PCB_FLUSH(Q)                                                                                  \
    if (TRACE) { *tp = Wacc; tp += T; }

#define PCB_GROUP(Q)                                                                                  \
    case 4 * (Q):     PCB_CELL(4 * (Q))                                                               \
    case 4 * (Q) + 1: PCB_CELL(4 * (Q) + 1)                                                           \
    case 4 * (Q) + 2: PCB_CELL(4 * (Q) + 2)                                                           \
    case 4 * (Q) + 3: PCB_CELL(4 * (Q) + 3)                                                           \
        PCB_FLUSH(Q)

}  // namespace

namespace {

// ------------------------------------------------------------------------------------------------
// k_align<TRACE>: TRACE=true  -> forward DP with flags + traceback + statistics -> pcb_result
//                 TRACE=false -> forward DP only -> best order key per task (long-read scan chunks)
// grid: persistent, blockDim = PCB_BLOCK; warp w of the grid takes units w, w + nwarps, ...
// dynamic smem: (PCB_BLOCK/32) * smax * PCB_RT words (adapter base pairs of each warp's unit)
// ------------------------------------------------------------------------------------------------
template <bool TRACE>
__global__ void __launch_bounds__(PCB_BLOCK, PCB_MIN_BLOCKS) k_align(const PcbLaunch L)
{
    extern __shared__ uint32_t s_ad_all[];
    constexpr int RT = PCB_RT;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t T = (uint64_t)gridDim.x * blockDim.x;                 // scratch stride (thread slots)
    const uint64_t slot = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t *const s_ad = s_ad_all + (size_t)warp * L.smax * RT;

    // per-thread scratch regions (all strided by T so that a warp's accesses coalesce)
    uint32_t *scr = L.scratch + slot;
    uint32_t *const tr_base = scr;                                         // [smax][mmax][RT/4]
    if (TRACE) scr += (uint64_t)L.smax * L.mmax * (RT / 4) * T;
    uint32_t *const bnd = scr;                                             // [mmax+1][2]
    scr += 2ull * (L.mmax + 1) * T;
    uint32_t *const dump = scr;                                            // [2][smax*RT][2]

    const uint32_t GO = pk2(L.sc.gap_open), GE = pk2(L.sc.gap_extend), GE1 = pk2(L.sc.gap_extend + 1);
    const uint32_t MATCHP = pk2(L.sc.match - L.sc.gap_open), XORC = MATCHP ^ pk2(L.sc.mismatch - L.sc.gap_open);
    const uint32_t NEGINF = pk2(PCB_NEG_INF);
    const int go = L.sc.gap_open, ge = L.sc.gap_extend;

    const int64_t nwarps = (int64_t)gridDim.x * (PCB_BLOCK / 32);
    for (int64_t unit = (int64_t)blockIdx.x * (PCB_BLOCK / 32) + warp; unit < L.src.n_units; unit += nwarps) {
        int32_t adapter = 0;
        HalfTask tk[2];
        fetch_task(L.src, unit, 2 * lane, adapter, tk[0]);
        fetch_task(L.src, unit, 2 * lane + 1, adapter, tk[1]);
        const int64_t aoff = L.ad.off[adapter];
        const int n = (int)(L.ad.off[adapter + 1] - aoff);
        if (n == 0) {
            if (TRACE) { for (int h = 0; h < 2; ++h) if (tk[h].out_idx >= 0) write_failed(L.out, tk[h].out_idx); }
            else { for (int h = 0; h < 2; ++h) if (tk[h].out_idx >= 0) L.scan_out[tk[h].out_idx] = (int)0x80000000; }
            continue;
        }
        const int S = (n + RT - 1) / RT;          // stripes
        const int r0 = S * RT - n;                // first used register row of stripe 0 (bottom aligned)
        __syncwarp();
        for (int i = lane; i < n; i += 32) s_ad[r0 + i] = (uint32_t)L.ad.codes[aoff + i] * 0x10001u;
        __syncwarp();

        int mloc = max(tk[0].len, tk[1].len);
        const int mw = __reduce_max_sync(0xffffffffu, mloc);
        // per half: does the segment end at the read's end (then its last column is SeqAn's final column)?
        const bool fin0 = tk[0].len > 0 && tk[0].col0 + tk[0].len == tk[0].m_true;
        const bool fin1 = tk[1].len > 0 && tk[1].col0 + tk[1].len == tk[1].m_true;
        const int jrow0 = fin0 ? tk[0].len - 1 : tk[0].len;     // last column whose last-row cell is row-class
        const int jrow1 = fin1 ? tk[1].len - 1 : tk[1].len;
        int best0 = (int)0x80000000, best1 = (int)0x80000000;
        // cell (n, 0): first tracked cell of the whole matrix, score 0 (dp_meta_info.h:211-213)
        if (tk[0].len > 0 && tk[0].col0 == 0 && tk[0].own0 == 0 && (tk[0].forced < 0 || tk[0].forced == 0)) best0 = make_key(0, 0, 0, 0);
        if (tk[1].len > 0 && tk[1].col0 == 0 && tk[1].own0 == 0 && (tk[1].forced < 0 || tk[1].forced == 0)) best1 = make_key(0, 0, 0, 0);
        const uint8_t *const seq0 = L.arena + tk[0].off;
        const uint8_t *const seq1 = L.arena + tk[1].off;
        const int lim0 = max(tk[0].len - 1, 0), lim1 = max(tk[1].len - 1, 0);

        for (int s = 0; s < S; ++s) {
            uint32_t Mo[RT], Hl[RT];
#pragma unroll
            for (int r = 0; r < RT; ++r) { Mo[r] = GO; Hl[r] = NEGINF; }   // column 0: M = 0, H = -inf
            const int r0s = (s == 0) ? r0 : 0;
            const uint32_t *const s_adp = s_ad + s * RT;
            const bool last_stripe = (s == S - 1);
            uint32_t prevTop = GO;
            for (int j = 1; j <= mw; ++j) {
                const uint32_t b0 = seq0[min(j - 1, lim0)], b1 = seq1[min(j - 1, lim1)];
                const uint32_t rp = b0 | (b1 << 16);
                uint32_t Mo_d, Mo_u, nMo_u, Vu, Wacc = 0;
                if (s == 0) { Mo_d = GO; Mo_u = GO; Vu = NEGINF; }          // row 0: M = 0, V = -inf
                else {
                    Mo_u = bnd[(2ull * j) * T]; Vu = bnd[(2ull * j + 1) * T];
                    Mo_d = prevTop; prevTop = Mo_u;
                }
                nMo_u = ~Mo_u;
                uint32_t *tp = tr_base + ((uint64_t)(s * L.mmax + (j - 1)) * (RT / 4) + (r0s >> 2)) * T;
                switch (r0s) {
                    PCB_GROUP(0) PCB_GROUP(1) PCB_GROUP(2) PCB_GROUP(3)
                    PCB_GROUP(4) PCB_GROUP(5) PCB_GROUP(6) PCB_GROUP(7)
                }
                if (!last_stripe) {
                    bnd[(2ull * j) * T] = Mo_u; bnd[(2ull * j + 1) * T] = Vu;
                } else {
                    // last-row cell of this column: scout candidate (both halves)
                    const uint32_t botM = vadd(Mo[RT - 1], pk2(-go));
                    const uint32_t xv = botM ^ Vu, xh = botM ^ Hl[RT - 1];
                    {
                        const bool ok = (j <= jrow0) && (j > tk[0].own0) && (tk[0].forced < 0 || tk[0].forced == j);
                        const int key = make_key(lo16(botM), j, (xv & 0xFFFFu) == 0, (xh & 0xFFFFu) == 0);
                        if (ok) best0 = max(best0, key);
                    }
                    {
                        const bool ok = (j <= jrow1) && (j > tk[1].own0) && (tk[1].forced < 0 || tk[1].forced == j);
                        const int key = make_key(hi16(botM), j, (xv >> 16) == 0, (xh >> 16) == 0);
                        if (ok) best1 = max(best1, key);
                    }
                }
                // final column of a half: keep the whole column (M+go and H of every row) for the post-pass
                const bool dmp0 = fin0 && j == tk[0].len, dmp1 = fin1 && j == tk[1].len;
                if (dmp0 || dmp1) {
                    uint32_t *p0 = dump + 2ull * (s * RT) * T, *p1 = p0 + 2ull * L.smax * RT * T;
                    if (!dmp0) p0 = p1;            // both areas get the same column when both halves end here
                    if (!dmp1) p1 = p0;
                    asm volatile("" : "+l"(p0), "+l"(p1));      // keep the address arithmetic inside this rare branch
#pragma unroll
                    for (int r = 0; r < RT; ++r) {
                        p0[0] = Mo[r]; p0[T] = Hl[r]; p1[0] = Mo[r]; p1[T] = Hl[r];
                        p0 += 2 * T; p1 += 2 * T;
                    }
                }
            }
        }

        // ---- post-pass: cells of the final column (adapter overhanging the read end), rows 1..n in order.
        //      V is rebuilt from the dumped column with the same recurrence the forward pass used.
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const bool fin = h ? fin1 : fin0;
            if (!fin) continue;
            const uint32_t *const dp = dump + (h ? 2ull * L.smax * RT * T : 0ull);
            int best = h ? best1 : best0;
            int Vp = PCB_NEG_INF, Mp = 0;                       // row 0: V = -inf, M = 0
            for (int i = 1; i <= n; ++i) {
                const int a = r0 + i - 1;
                const int M = half16(dp[(2ull * a) * T], h) - go;
                const int H = half16(dp[(2ull * a + 1) * T], h);
                const int V = max(Vp + ge, Mp + go);
                const int ord = tk[h].len + i;
                if (tk[h].forced < 0 || tk[h].forced == ord) best = max(best, make_key(M, ord, V == M, H == M));
                Vp = V; Mp = M;
            }
            if (h) best1 = best; else best0 = best;
        }

        if (!TRACE) {
            if (tk[0].out_idx >= 0) L.scan_out[tk[0].out_idx] = tk[0].len > 0 ? best0 : (int)0x80000000;
            if (tk[1].out_idx >= 0) L.scan_out[tk[1].out_idx] = tk[1].len > 0 ? best1 : (int)0x80000000;
            continue;
        }

        // ---- traceback + statistics, one half after the other (dp_traceback_impl.h:379-451, :498-545)
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            const HalfTask &t = tk[h];
            if (t.out_idx < 0) continue;
            if (t.len <= 0) { write_failed(L.out, t.out_idx); continue; }
            const int best = h ? best1 : best0;
            if (best == (int)0x80000000) {          // forced cell not found: report through the error flag
                atomicOr(L.err_flag, 2);
                write_failed(L.out, t.out_idx);
                continue;
            }
            const int score = best >> 16;
            const int ord = 0x3FFF - ((best >> 2) & 0x3FFF);
            const int eqV = (best >> 1) & 1, eqH = best & 1;
            int ci, cj;
            if (ord <= t.len) { ci = n; cj = ord; } else { ci = ord - t.len; cj = t.len; }
            const int end_i = ci, end_j = cj;
            // _correctTraceValue (affine only): V == M beats H == M beats the recorded direction
            int state = 0;                                        // 0 = M, 1 = V, 2 = H
            if (!L.sc.linear) state = eqV ? 1 : (eqH ? 2 : 0);
            PathAcc acc; path_acc_init(acc);
            const uint8_t *const seq = L.arena + t.off;
            while (ci > 0 && cj > 0) {
                const int a = r0 + ci - 1;
                const int s = a / RT, r = a % RT;
                const uint32_t w = tr_base[((uint64_t)(s * L.mmax + (cj - 1)) * (RT / 4) + (r >> 2)) * T];
                const int sh = 8 * h + 2 * (r & 3);
                const int fh = (w >> sh) & 1, fg = (w >> (sh + 1)) & 1;
                const int fv = (w >> (16 + sh)) & 1, fd = (w >> (17 + sh)) & 1;
                if (state == 0) {
                    if (fd) {
                        const int eq = (seq[cj - 1] == (uint8_t)(s_ad[a] & 0xFFu));
                        path_acc_step(acc, 0, eq);
                        --ci; --cj;
                    } else {
                        state = fg ? 1 : 2;
                    }
                } else if (state == 1) {
                    path_acc_step(acc, 1, 0);
                    --ci;
                    state = (fv || L.sc.linear) ? 0 : 1;
                } else {
                    path_acc_step(acc, 2, 0);
                    --cj;
                    state = (fh || L.sc.linear) ? 0 : 2;
                }
            }
            if (cj == 0 && ci > 0 && t.col0 > 0) atomicOr(L.err_flag, 1);   // window too narrow for this path
            PathStats st;
            path_finalize(acc, (int64_t)t.col0 + cj, ci, (int64_t)t.m_true - (t.col0 + end_j), n - end_i,
                          t.m_true, n, st);
            pcb_result res;
            res.read_start = st.read_start; res.read_end = st.read_end;
            res.adapter_start = st.adapter_start; res.adapter_end = st.adapter_end;
            res.raw_score = score; res.matches = st.matches;
            res.aligned_len = st.aligned_len; res.full_len = st.full_len;
            L.out[t.out_idx] = res;
            if (L.aux) L.aux[t.out_idx] = (uint8_t)st.end_match;
        }
    }
}

}  // namespace

namespace {

// One thread per long read.  Chunk q owns columns (q*chunk, min(len, (q+1)*chunk)] and starts `warm` columns early;
// its key holds the best owned cell as (score << 16) | ((0x3FFF - local order) << 2).  Global visiting order
// (dp_scout.h:167-180, dp_meta_info.h:205-216): the column for last-row cells, len + row for final-column cells.
__global__ void k_merge_scan(const PcbLong *__restrict__ longs, int64_t n_long, const int32_t *__restrict__ keys,
                             PcbTask *__restrict__ wins, int32_t *err_flag)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_long) return;
    const PcbLong lt = longs[t];
    int best_score = 0; int64_t best_ord = 0; bool have = false;
    for (int q = 0; q < lt.nchunks; ++q) {
        const int32_t key = keys[lt.chunk0 + q];
        if (key == (int32_t)0x80000000) continue;
        const int64_t c0 = (int64_t)q * lt.chunk;
        const int64_t col0 = c0 > lt.warm ? c0 - lt.warm : 0;
        const int64_t cend = c0 + lt.chunk < lt.len ? c0 + lt.chunk : lt.len;
        const int64_t chlen = cend - col0;
        const int score = key >> 16;
        const int ord_l = 0x3FFF - ((key >> 2) & 0x3FFF);
        const int64_t ord_g = ord_l <= chlen ? col0 + ord_l : (int64_t)lt.len + (ord_l - chlen);
        if (!have || score > best_score || (score == best_score && ord_g < best_ord)) { best_score = score; best_ord = ord_g; have = true; }
    }
    PcbTask d;
    d.own0 = 0; d.out_idx = lt.out_idx; d.m_true = lt.len;
    if (!have) {                       // cannot happen (cell (n, 0) always scores 0); reported, never silent
        atomicOr(err_flag, 8);
        d.off = (uint64_t)lt.off; d.len = 1; d.col0 = 0; d.forced_ord = 0x3FFF;
    } else {
        const bool rowclass = best_ord < lt.len;
        const int64_t cj = rowclass ? best_ord : lt.len;
        const int64_t ws = cj > (int64_t)lt.warm + lt.span ? cj - ((int64_t)lt.warm + lt.span) : 0;
        d.col0 = (int32_t)ws; d.len = (int32_t)(cj - ws > 1 ? cj - ws : 1); d.off = (uint64_t)(lt.off + ws);
        d.forced_ord = rowclass ? (int32_t)(cj - ws) : (int32_t)(d.len + (best_ord - lt.len));
    }
    wins[lt.win_slot] = d;
}

// Planner, two passes over the caller's task list (one thread per task): FILL=false counts items per bucket,
// FILL=true writes them at base[bucket] + (atomic cursor).  Which slot inside a bucket a task gets varies from run
// to run; results do not (tasks are independent and carry their own result slot).
__device__ __forceinline__ void plan_classify(const PcbPlanArgs &a, int32_t len, const PcbAdInfo &x, bool &is_long,
                                              int &nchunks, int &chunk_r, int &lb)
{
    is_long = x.chunked && len > x.direct_max && x.n > 0;
    if (!is_long) { nchunks = 0; chunk_r = 0; lb = min(len >> 4, PCB_LB - 1); return; }
    nchunks = (len + a.chunk - 1) / a.chunk;
    chunk_r = ((len + nchunks - 1) / nchunks + 3) & ~3;          // near-equal chunks: a warp's lanes stay busy together
    nchunks = (len + chunk_r - 1) / chunk_r;
    lb = min((chunk_r + x.warm) >> 7, PCB_LB - 1);
}

template <bool FILL>
__global__ void k_plan(const PcbPlanArgs a)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n_tasks) return;
    const int32_t len = a.seg_len[t], ad = a.seg_ad[t];
    const int64_t off = a.seg_off[t];
    if (off + len <= a.end_lo || off + len > a.end_hi) return;      // another piece's task
    const PcbAdInfo x = a.ad[ad];
    bool is_long; int nchunks, chunk_r, lb;
    plan_classify(a, len, x, is_long, nchunks, chunk_r, lb);
    const int64_t per_kind = (int64_t)a.n_adapters * PCB_LB;
    if (!is_long) {
        const int64_t bkt = (int64_t)ad * PCB_LB + lb;
        if (!FILL) { atomicAdd(a.count + bkt, 1); atomicMax(a.mmax + x.cls, len); return; }
        const int64_t slot = a.base[bkt] + atomicAdd(a.cursor + bkt, 1);
        PcbTask d; d.off = (uint64_t)off; d.len = len; d.col0 = 0; d.m_true = len; d.forced_ord = -1;
        d.out_idx = (int32_t)t; d.own0 = 0;
        a.tasks[slot] = d;
        return;
    }
    const int64_t bs = per_kind + (int64_t)ad * PCB_LB + lb, bw = 2 * per_kind + (int64_t)ad * PCB_LB;
    if (!FILL) {
        atomicAdd(a.count + bs, nchunks); atomicAdd(a.count + bw, 1);
        atomicMax(a.mmax + (PCF_NCLS + 1) + x.cls, min(len, chunk_r + x.warm));
        return;
    }
    const int64_t s0 = a.base[bs] + atomicAdd(a.cursor + bs, nchunks);
    const int64_t w = a.base[bw] + atomicAdd(a.cursor + bw, 1);
    for (int q = 0; q < nchunks; ++q) {
        const int64_t c0 = (int64_t)q * chunk_r;
        const int64_t col0 = c0 > x.warm ? c0 - x.warm : 0;
        const int64_t cend = c0 + chunk_r < len ? c0 + chunk_r : len;
        PcbTask d; d.off = (uint64_t)(off + col0); d.len = (int32_t)(cend - col0); d.col0 = (int32_t)col0;
        d.m_true = len; d.forced_ord = -1; d.out_idx = (int32_t)(s0 + q - a.scan_task0); d.own0 = (int32_t)(c0 - col0);
        a.tasks[s0 + q] = d;
    }
    PcbLong lt; lt.off = off; lt.len = len; lt.warm = x.warm; lt.span = x.span; lt.chunk = chunk_r;
    lt.chunk0 = (int32_t)(s0 - a.scan_task0); lt.nchunks = nchunks; lt.out_idx = (int32_t)t;
    lt.win_slot = (int32_t)(w - a.win_task0);
    a.longs[w - a.win_task0] = lt;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host-side launchers
// ------------------------------------------------------------------------------------------------
cudaError_t pcb_launch_plan(const PcbPlanArgs &a, bool fill, cudaStream_t st)
{
    if (a.n_tasks == 0) return cudaSuccess;
    const unsigned grid = (unsigned)((a.n_tasks + 255) / 256);
    if (fill) k_plan<true><<<grid, 256, 0, st>>>(a); else k_plan<false><<<grid, 256, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t pcb_launch_merge(const PcbLong *longs, int64_t n_long, const int32_t *keys, PcbTask *wins, int32_t *err_flag,
                             cudaStream_t st)
{
    if (n_long == 0) return cudaSuccess;
    k_merge_scan<<<(unsigned)((n_long + 127) / 128), 128, 0, st>>>(longs, n_long, keys, wins, err_flag);
    return cudaGetLastError();
}

cudaError_t pcb_launch_gather_windows(const uint8_t *arena_t, const int64_t *roff, const int32_t *rlen, int64_t n, int end_size, int at_end,
                                      int stride, uint8_t *arena, int32_t *win_len, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    const int64_t threads = n * (stride / 4);
    k_gather_windows<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(arena_t, roff, rlen, n, end_size, at_end, stride, arena, win_len);
    return cudaGetLastError();
}

cudaError_t pcb_launch_encode(const char *ascii, uint8_t *codes, uint64_t n, cudaStream_t st)
{
    if (n == 0) return cudaSuccess;
    const uint64_t threads = (n + 15) / 16;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    k_encode<<<grid, 256, 0, st>>>(ascii, codes, n);
    return cudaGetLastError();
}

static size_t smem_bytes(const PcbLaunch &L) { return (size_t)(PCB_BLOCK / 32) * L.smax * PCB_RT * sizeof(uint32_t); }

cudaError_t pcb_launch_trace(const PcbLaunch &L, cudaStream_t st)
{
    if (L.src.n_units == 0) return cudaSuccess;
    k_align<true><<<L.grid, PCB_BLOCK, smem_bytes(L), st>>>(L);
    return cudaGetLastError();
}

cudaError_t pcb_launch_scan(const PcbLaunch &L, cudaStream_t st)
{
    if (L.src.n_units == 0) return cudaSuccess;
    k_align<false><<<L.grid, PCB_BLOCK, smem_bytes(L), st>>>(L);
    return cudaGetLastError();
}

int pcb_occupancy_grid(int which)
{
    int dev = 0, sms = 0, per = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (which == 0) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_align<true>, PCB_BLOCK, 4 * 4 * PCB_RT * 4);
    else            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_align<false>, PCB_BLOCK, 4 * 4 * PCB_RT * 4);
    if (per < 1) per = 1;
    return sms * per;
}
