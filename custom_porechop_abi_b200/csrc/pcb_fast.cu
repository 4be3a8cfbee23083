/*
 * pcb_fast.cu -- the fast sm_100a kernels of the adapter-alignment hot path (affine schemes, ACGT adapters of <= 72 bases).
 *
 * Same result contract as pcb_kernels.cu (reference: porechop_abi/src/adapter_align.cpp:11-31 = SeqAn 2.4.0 globalAlignment
 * with all four end gaps free, Gotoh recursion seqan/align/dp_formula_affine.h:65-124, single trace, then
 * porechop_abi/src/alignment.cpp:5-110), but organised around what the B200 integer pipes cost (measured with
 * scripts/microbench/pipes.cu: every packed 16x2 / DPX / PRMT / LOP3 / IMAD instruction is 2 clk per warp and sub-partition,
 * 32-bit adds about 1.4 clk, and FMA-pipe work hides only a little under ALU work), i.e. around the instruction count per cell:
 *
 *   1. anti-diagonal bias.  Every DP value is stored as  x + b*(i+j) + B0  with b = -gapOpen.  All four candidates of a
 *      cell share the bias, so every comparison (and therefore every tie-break) is unchanged, but
 *        - M + gapOpen needs no instruction (gapOpen + b == 0),
 *        - all stored values and the substitution scores (s + 2b) are non-negative, so the diagonal add is a plain
 *          32-bit add on the packed pair (no carry can cross the halves).
 *      Score-only cell for two tasks: LDS (substitution scores of the two read bases against the adapter base, from a
 *      per-warp table in shared memory), 32-bit add, 2 x VIADDMNMX.S16x2 (H and V), VIMNMX3.S16x2 (M): 4 integer
 *      instructions; the lean column loop of the built library runs at 5.75 instructions per cell pair.
 *   2. two passes, two kernels, instead of a trace matrix.  k_fast<1> (pass 1) is score-only: it finds SeqAn's end cell
 *      (dp_scout.h:167-180) and leaves the DP column of every C-th column in HBM (one 2*NR-word record per thread and
 *      checkpoint, streaming stores).  k_rounds (pass 2) recomputes, with trace flags, only the C-column blocks a traceback
 *      actually walks through (1.8 of them on average) -- and of those only the columns and rows the walk can reach, 64
 *      tracebacks of a kind per warp (they are sorted first) --, keeps the flags in shared memory and walks them.  Values are
 *      15-bit non-negative, so "a >= b" is bit 15 of (a + 0x8000 - b): one IADD3 per flag, extracted four at a time with
 *      PRMT in sign-replicate mode.  Two kernels because the passes want different things: pass 1 is a register-bound
 *      integer stream with no memory stalls, the rounds are short latency-bound phases (checkpoint fetch, flag walk) that
 *      need flag memory and resident warps; fused into one kernel (round 1) each held the other back.
 *   3. the scan kernel for long reads (middle-adapter search) is pass 1 without checkpoints (k_fast<0>), with the bias
 *      re-based every 1024 columns.
 *   4. no copies, no waits.  The column body is straight-line code over NR rows (template parameter: 24 = the length of
 *      almost every Porechop adapter and barcode, exact; 24 / 32 / 40 / 52 = any shorter-or-equal adapter, bottom-aligned,
 *      72 for the 63- and 68-base native full-sequence barcode sets -- the rows above the adapter being exact copies of DP
 *      row 0), so ptxas updates the DP registers in place; read bases
 *      are fetched four columns at a time, one group ahead of their use, so that no column waits on a global load.
 *
 * One thread = two tasks (the halves of each 32-bit register), one warp = 64 tasks sharing an adapter, as in pcb_kernels.cu.
 * Sequence codes in the arenas are Dna5 code * 8 (0, 8, 16, 24, 32 = N), so that a code is directly a byte shift.
 */
#include "pcb_kernels.h"
#include "align_stats.h"

#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>

namespace {

// CTAs per SM the register allocation is held to.  Measured on B200: the score-only kernels issue integer instructions at the
// pipes' limit with 2-3 CTAs per SM already; a larger register budget (no spills, fewer register moves) beats more warps.
#ifndef PCF_MINB_SCAN
#define PCF_MINB_SCAN 2
#endif
#ifndef PCF_MINB_P1
#define PCF_MINB_P1(NR) ((NR) <= 24 ? 4 : ((NR) <= 32 ? 3 : 2))
#endif
#ifndef PCF_MINB_P2
#define PCF_MINB_P2(NR) ((NR) <= 24 ? 4 : 2)
#endif
#define PCF_MINB(MODE, NR) ((MODE) ? PCF_MINB_P1(NR) : PCF_MINB_SCAN)
#ifndef PCF_UNROLL
#define PCF_UNROLL 2
#endif
// -DPCF_CHECKED: every index the trace kernels compute into scratch, shared memory and the result array is range-checked
// (device error flag bit 6); the GPU tests are run once against such a build (scripts/gpu_checked.sh -> profiles/)
#ifdef PCF_CHECKED
#define PCF_CHECK(L, cond) do { if (!(cond)) atomicOr((L).err_flag, 64); } while (0)
#else
#define PCF_CHECK(L, cond) do { } while (0)
#endif
constexpr int UNR = PCF_UNROLL;
#ifndef PCF_UNROLL2
#define PCF_UNROLL2 1
#endif
constexpr int UNR2 = PCF_UNROLL2;
constexpr int B0 = 64;                // base bias: keeps every stored value > 0
constexpr int OFF = 2048;             // offset of un-biased scores in the best-cell comparison (true score + OFF > 0)
constexpr int REBASE = 1024;          // scan: columns between two re-basings of the anti-diagonal bias
constexpr uint32_t GUARD = 0x80008000u;

__device__ __forceinline__ uint32_t pk2(int v) { return ((uint32_t)v & 0xFFFFu) * 0x10001u; }
__device__ __forceinline__ uint32_t pk(int lo, int hi) { return ((uint32_t)lo & 0xFFFFu) | ((uint32_t)hi << 16); }
__device__ __forceinline__ int lo16(uint32_t x) { return (int)(short)(x & 0xFFFFu); }
__device__ __forceinline__ int hi16(uint32_t x) { return (int)x >> 16; }
__device__ __forceinline__ int half16(uint32_t x, int h) { return h ? hi16(x) : lo16(x); }
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t s)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(s));
    return d;
}
// bytes: sign(a.lo), sign(a.hi), sign(b.lo), sign(b.hi), each replicated to 0x00 / 0xFF
__device__ __forceinline__ uint32_t sign4(uint32_t a, uint32_t b)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, 0xfdb9;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
// per-half max(a, b) with keep_x = (a_x >= b_x).  Same PTX pattern as the toolkit's __vibmax_s16x2 (ptxas turns it into one
// VIMNMX.S16x2 with two predicate outputs), but with an early-clobber result: the header's version lets the result share
// a register with `a`, and then compares the result with itself.
__device__ __forceinline__ uint32_t vmax_keep(uint32_t a, uint32_t b, bool &keep_hi, bool &keep_lo)
{
    uint32_t val, ph, pl;
    asm("{.reg .pred pu, pv; \n\t"
        ".reg .s16 rs0, rs1, rs2, rs3; \n\t"
        "max.s16x2 %0, %3, %4; \n\t"
        "mov.b32 {rs0, rs1}, %0; \n\t"
        "mov.b32 {rs2, rs3}, %3; \n\t"
        "setp.eq.s16 pv, rs0, rs2; \n\t"
        "setp.eq.s16 pu, rs1, rs3; \n\t"
        "selp.b32 %1, 1, 0, pu; \n\t"
        "selp.b32 %2, 1, 0, pv;} \n\t"
        : "=&r"(val), "=r"(ph), "=r"(pl) : "r"(a), "r"(b));
    keep_hi = ph != 0; keep_lo = pl != 0;
    return val;
}
// a + GUARD - b on the packed pair: bit 15 / 31 = (a >= b) per half, for 15-bit non-negative a, b
__device__ __forceinline__ uint32_t cmpge(uint32_t a, uint32_t b) { return a + GUARD - b; }
// 256-bit streaming accesses (sm_100a: STG / LDG.E.EF.ENL2.256): a warp moves 1 KB per instruction
__device__ __forceinline__ void st256(uint32_t *p, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7)
{
    asm volatile("st.global.cs.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" :: "l"(p), "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7) : "memory");
}
__device__ __forceinline__ void ld256(const uint32_t *p, uint32_t (&v)[8])
{
    asm volatile("ld.global.cs.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];" : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]) : "l"(p));
}

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
    pcb_result r;
    r.read_start = -1; r.read_end = 0; r.adapter_start = -1; r.adapter_end = 0;
    r.raw_score = (int32_t)0x80000000; r.matches = 0; r.aligned_len = 0; r.full_len = 0;
    out[idx] = r;
}


// ------------------------------------------------------------------------------------------------
// Score table of a warp's adapter, in shared memory: tbl[r * 25 + p] = packed (score(b0, adapter[r]) + 2b,
// score(b1, adapter[r]) + 2b) for the pair p = 5 * code(b0) + code(b1) of read bases the two tasks of a thread see in
// a column.  A cell's substitution score is then ONE shared-memory load at [per-column pointer + constant row offset]:
// no ALU instruction (the 25 words of a row lie in 25 different banks: conflict-free for any mix of lanes).
// ------------------------------------------------------------------------------------------------
constexpr int TP = 25;

// One DP column, score only, for the two tasks of this thread.  Mo[r] / Hl[r]: M and H of the left neighbour (biased),
// updated in place.  rowTp / rowT: biased M(0, j-1) / M(0, j).  tp = tbl + p of this column.  The diagonal term of the
// NEXT row (old Mo[r] + score) is formed before Mo[r] is overwritten.  Per row: LDS, IADD, VIADDMNMX (H),
// VIADDMNMX (V), VIMNMX3 (M).  Recurrence: seqan/align/dp_formula_affine.h:65-124.
// Ml / Vl / Hlast: M, V, H of the adapter's last row = register row NR-1 (adapters are bottom-aligned, see k_fast).
template <int NR, bool EXACT>
__device__ __forceinline__ void col_score(uint32_t (&Mo)[NR], uint32_t (&Hl)[NR], const uint32_t *tp,
                                          uint32_t rowTp, uint32_t rowT, uint32_t GEB, uint32_t &Ml)
{
    uint32_t Dn = rowTp + tp[0];
    uint32_t Vu = rowT;                              // V(1, j) = M(0, j) + gapOpen  (row 0: V = -inf)
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const uint32_t Hn = __viaddmax_s16x2(Hl[r], GEB, Mo[r]);
        if (r > 0) Vu = __viaddmax_s16x2(Vu, GEB, Mo[r - 1]);
        const uint32_t D = Dn;
        if (r < NR - 1) Dn = Mo[r] + tp[(r + 1) * TP];
        Mo[r] = __vimax3_s16x2(D, Vu, Hn); Hl[r] = Hn;
    }
    Ml = Mo[NR - 1];
}

// The same column with trace flags (pass 2).  Flag bits of row r, half h, in flag word r >> 2 with q = r & 3:
//      bit 8h+2q   : H extended (H_left + ge >= M_left + go)     bit 8h+2q+1    : V >= H
//      bit 16+8h+2q: V extended                                  bit 16+8h+2q+1 : diagonal >= max(V, H)
// (tie rules of dp_formula_affine.h:65-124 / dp_formula.h:147-163: the left operand of each _maxScore wins ties).
// "x >= y" is bit 15 of x + GUARD - y (one IADD3) for 15-bit non-negative x, y.  fp: this thread's flag words of the
// column, stride PCB_BLOCK.  Only register rows < rlim (a multiple of 4, the same for the whole warp) are computed: a
// traceback that stands in row i never looks at a row below i.
template <int NR, bool EXACT>
__device__ __forceinline__ void col_flags(uint32_t (&Mo)[NR], uint32_t (&Hl)[NR], const uint32_t *tp,
                                          uint32_t rowTp, uint32_t rowT, uint32_t GEB, uint32_t *fp, const int rlim,
                                          uint32_t &Ml, uint32_t &Vl, uint32_t &Hlast)
{
    // the substitution scores of the column come from shared memory PF rows ahead of their use
#ifndef PCF_PF
#define PCF_PF 8
#endif
    constexpr int PF = PCF_PF;
    uint32_t tv[PF];
#pragma unroll
    for (int d = 0; d < PF; ++d) tv[d] = tp[d * TP];
    uint32_t Dn = rowTp + tv[0];
    uint32_t Vu = 0, Wacc = 0;
    const uint32_t GEBG = GEB + GUARD;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        if ((r & 3) == 0 && r > 0 && r >= rlim) break;
        const uint32_t D = Dn;
        if (r + PF < NR) tv[r % PF] = tp[(r + PF) * TP];
        const uint32_t Mup = (r == 0) ? rowT : Mo[(r > 0) ? r - 1 : 0];
        const uint32_t ch = Hl[r] + GEBG - Mo[r], cv = Vu + GEBG - Mup;
        const uint32_t Hn = __viaddmax_s16x2(Hl[r], GEB, Mo[r]), Vn = __viaddmax_s16x2(Vu, GEB, Mup);
        const uint32_t G = __vmaxs2(Vn, Hn);
        const uint32_t W = (sign4(ch, cv) & 0x55555555u) | (sign4(cmpge(Vn, Hn), cmpge(D, G)) & 0xAAAAAAAAu);
        if ((r & 3) == 0) Wacc = W;
        else if ((r & 3) == 1) Wacc = (Wacc & 0x03030303u) | (W & 0xFCFCFCFCu);
        else if ((r & 3) == 2) Wacc = (Wacc & 0x0F0F0F0Fu) | (W & 0xF0F0F0F0u);
        else Wacc = (Wacc & 0x3F3F3F3Fu) | (W & 0xC0C0C0C0u);
        if ((r & 3) == 3 || r == NR - 1) fp[(r >> 2) * PCB_BLOCK] = Wacc;
        if (r < NR - 1) Dn = Mo[r] + tv[(r + 1) % PF];
        Vu = Vn; Mo[r] = __vmaxs2(D, G); Hl[r] = Hn;
    }
    Ml = Mo[NR - 1]; Vl = Vu; Hlast = Hl[NR - 1];
}

// Four consecutive sequence bytes starting at seq[pos] (pos clamped to lim) come from two aligned words; the funnel shift
// that joins them is done where the bytes are used, so that the loads can be issued a group of columns ahead.
struct Raw4 { uint32_t w0, w1, sh; };
__device__ __forceinline__ Raw4 load4(const uint8_t *seq, int pos, int lim)
{
    const uintptr_t a = (uintptr_t)(seq + min(pos, lim));
    const uint32_t *p = reinterpret_cast<const uint32_t *>(a & ~(uintptr_t)3);
    Raw4 r; r.w0 = p[0]; r.w1 = p[1]; r.sh = (uint32_t)(a & 3) * 8u;
    return r;
}
__device__ __forceinline__ uint32_t join4(const Raw4 &r) { return __funnelshift_r(r.w0, r.w1, r.sh); }
// offset (in words) into a row of the score table for the two code bytes (Dna5 code * 8) of a column
__device__ __forceinline__ uint32_t pair_index(uint32_t c8_0, uint32_t c8_1) { return (c8_0 * 5u + c8_1) >> 3; }

__device__ __forceinline__ uint64_t ckw_of(int mmax, int C, int NR) { return (uint64_t)((mmax + C - 1) / C) * (uint64_t)((NR + 7) / 8) * 8ull; }   // = pcf_ck_words (pcb_kernels.h)

// ------------------------------------------------------------------------------------------------
// k_fast<MODE, NR, EXACT, C>: the score-only pass
//   MODE 0 -> scan: order key per task (chunks of long reads)
//   MODE 1 -> pass 1 of the trace: the DP column of every C-th column goes to L.scratch (checkpoints), the end cell of
//             every task (SeqAn's scout, dp_scout.h:167-180) to L.entries; k_rounds takes it from there
//   NR / EXACT  -> adapter rows in registers; EXACT: every adapter of the launch has exactly NR bases
//   C           -> checkpoint distance = width of a flag block (columns; a multiple of 4, >= 8)
// grid: persistent, blockDim = PCB_BLOCK, units [L.unit0, L.unit1).  Cross-product launches are walked adapter-major in
// contiguous per-warp ranges (a warp re-builds its score table only when the adapter changes); explicit task lists are
// walked strided.
// dynamic smem: [W][NR*25] score tables, [W][NR] adapter codes, then 2*NR words per thread (strided by PCB_BLOCK) for the
//               final-column dump.
// Entry of a task (two words): x = end row | end column << 7 | start state << 17 | 1 << 31 (0: no entry, the failed record
// is already written), y = raw score + OFF.  Start state 4 = "decide from V == M / H == M of the end cell".
// ------------------------------------------------------------------------------------------------
template <int MODE, int NR, bool EXACT, int C>
__global__ void __launch_bounds__(PCB_BLOCK, PCF_MINB(MODE, NR)) k_fast(const PcbLaunch L)
{
    static_assert(C % 4 == 0 && C >= 8, "flag block width");
    constexpr bool TRACE = MODE != 0;
    extern __shared__ uint32_t smem[];
    constexpr int W = PCB_BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *const tbl = smem + warp * (NR * TP);
    uint32_t *const s_code = smem + W * (NR * TP) + warp * NR;
    uint32_t *const s_fl = smem + W * (NR * TP + NR) + threadIdx.x;
    const uint64_t ckw = ckw_of(L.mmax, C, NR);

    const int b = -L.sc.gap_open;
    const uint32_t GEB = pk2(L.sc.gap_extend + b), Bp = pk2(b);
    const int s_match = L.sc.match + 2 * b, s_mis = L.sc.mismatch + 2 * b;
    const int geb = L.sc.gap_extend + b;

    const int64_t nwarps = (int64_t)gridDim.x * W, gw = (int64_t)blockIdx.x * W + warp;
    const bool cross = L.src.units == nullptr;
    const int64_t groups = cross ? (L.src.n_windows + 63) / 64 : 0;
    const int64_t nu = L.unit1 - L.unit0;
    int64_t u_begin, u_end, u_step;
    if (cross) { u_begin = L.unit0 + gw * nu / nwarps; u_end = L.unit0 + (gw + 1) * nu / nwarps; u_step = 1; }
    else       { u_begin = L.unit0 + gw; u_end = L.unit1; u_step = nwarps; }
    int32_t tbl_adapter = -1;

    for (int64_t unit = u_begin; unit < u_end; unit += u_step) {
        int32_t adapter = 0;
        HalfTask tk[2];
        if (cross) {                                     // adapter-major: unit = k * groups + g
            const int64_t k = unit / groups, g = unit - k * groups;
            adapter = L.src.ids[k];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                HalfTask &t = tk[h];
                t.off = 0; t.len = 0; t.col0 = 0; t.m_true = 0; t.forced = -1; t.own0 = 0; t.out_idx = -1;
                const int64_t w = g * 64 + 2 * lane + h;
                if (w < L.src.n_windows) {
                    t.off = (uint64_t)w * (uint64_t)L.src.stride;
                    t.len = L.src.win_len[w]; t.m_true = t.len;
                    t.out_idx = w * L.src.n_ids_total + (L.src.id_pos ? L.src.id_pos[k] : (int32_t)k);
                }
            }
        } else {
            fetch_task(L.src, unit, 2 * lane, adapter, tk[0]);
            fetch_task(L.src, unit, 2 * lane + 1, adapter, tk[1]);
        }
        uint4 *const entp = TRACE ? L.entries + (unit - L.unit0) * 32 + lane : nullptr;
        const int64_t aoff = L.ad.off[adapter];
        const int n = (int)(L.ad.off[adapter + 1] - aoff);
        if (n <= 0 || n > NR || (EXACT && n != NR)) {          // the host never sends these here; fail loudly if it does
            atomicOr(L.err_flag, 4);
            for (int h = 0; h < 2; ++h) if (tk[h].out_idx >= 0) {
                if (TRACE) write_failed(L.out, tk[h].out_idx); else L.scan_out[tk[h].out_idx] = (int)0x80000000;
            }
            if (TRACE) *entp = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
        // The adapter is bottom-aligned in the NR register rows: adapter row i lives in register row r0 + i - 1 and the
        // last adapter row is always row NR-1.  The r0 rows above it score 0 against every base, which makes each of them
        // an exact copy of DP row 0 (M = 0 by the diagonal, H and V never above gapOpen), so adapter row 1 sees the same
        // neighbours as under row 0 itself and no row of the column loop needs a bounds test.
        const int r0 = EXACT ? 0 : NR - n;
        if (adapter != tbl_adapter) {
            __syncwarp();
            for (int i = lane; i < NR; i += 32) s_code[i] = i >= r0 ? (uint32_t)(L.ad.codes[aoff + i - r0] / PCB_CODE_SCALE) : 0u;
            __syncwarp();
            for (int e = lane; e < NR * TP; e += 32) {
                const int r = e / TP, p = e - r * TP;
                const uint32_t a = s_code[r], c0 = (uint32_t)(p / 5), c1 = (uint32_t)(p % 5);
                tbl[e] = r >= r0 ? pk(c0 == a ? s_match : s_mis, c1 == a ? s_match : s_mis) : pk(2 * b, 2 * b);
            }
            __syncwarp();
            tbl_adapter = adapter;
        }

        const int len0 = tk[0].len, len1 = tk[1].len;
        const int mw = __reduce_max_sync(0xffffffffu, max(len0, len1));
        const bool fin0 = len0 > 0 && tk[0].col0 + len0 == tk[0].m_true;
        const bool fin1 = len1 > 0 && tk[1].col0 + len1 == tk[1].m_true;
        // row-class (last adapter row) candidate columns [lo, hi] per half
        int lo0 = tk[0].own0 + 1, hi0 = fin0 ? len0 - 1 : len0;
        int lo1 = tk[1].own0 + 1, hi1 = fin1 ? len1 - 1 : len1;
        if (tk[0].forced >= 0) { if (tk[0].forced <= hi0) { lo0 = max(lo0, tk[0].forced); hi0 = tk[0].forced; } else hi0 = -1; }
        if (tk[1].forced >= 0) { if (tk[1].forced <= hi1) { lo1 = max(lo1, tk[1].forced); hi1 = tk[1].forced; } else hi1 = -1; }
        // "lean" columns [ulo, uhi]: every task of the warp takes its row-class candidate and none is at its final column
        const bool act0 = tk[0].out_idx >= 0 && len0 > 0, act1 = tk[1].out_idx >= 0 && len1 > 0;
        const int ulo = __reduce_max_sync(0xffffffffu, max(act0 ? lo0 : 0, act1 ? lo1 : 0));
        const int ufin = __reduce_min_sync(0xffffffffu, min(fin0 ? len0 : 0x7fffffff, fin1 ? len1 : 0x7fffffff));
        const int uhi = min(__reduce_min_sync(0xffffffffu, min(act0 ? hi0 : 0x7fffffff, act1 ? hi1 : 0x7fffffff)), ufin - 1);
        // cell (n, 0): first tracked cell of the matrix, score 0 (dp_meta_info.h:211-213)
        const bool z0 = len0 > 0 && tk[0].col0 == 0 && tk[0].own0 == 0 && (tk[0].forced < 0 || tk[0].forced == 0);
        const bool z1 = len1 > 0 && tk[1].col0 == 0 && tk[1].own0 == 0 && (tk[1].forced < 0 || tk[1].forced == 0);
        uint32_t bestS = pk(z0 ? OFF : 0, z1 ? OFF : 0);        // un-biased score + OFF per half; 0 = none yet
        int ord0 = 0, ord1 = 0;
        int finS0 = 0, finS1 = 0, finI0 = 0, finI1 = 0, finE0 = 0, finE1 = 0;   // best final-column cell: score+OFF, row, eq bits
        const uint8_t *const seq0 = L.arena + tk[0].off;
        const uint8_t *const seq1 = L.arena + tk[1].off;
        const int lim0 = max(len0 - 1, 0), lim1 = max(len1 - 1, 0);

        uint32_t Mo[NR], Hl[NR];
#pragma unroll
        for (int r = 0; r < NR; ++r) { Mo[r] = pk2(B0 + b * (r + 1)); Hl[r] = 0; }   // column 0: M = 0, H = -inf
        uint32_t rowT = pk2(B0);                                 // biased M(0, j) = bias of row 0 at the current column
        const uint32_t candK = pk2(OFF - b * NR);                // candidate = M(n, j) + candK - rowT = score + OFF
        // this thread's part of the unit's checkpoints: [unit - unit0][k - 1][(NR + 7) / 8][lane][8 words = 8 rows, 16 bits per task]
        uint32_t *const ckp = TRACE ? L.scratch + (uint64_t)(unit - L.unit0) * 32u * ckw + (uint64_t)lane * 8u : nullptr;
        uint32_t ckoff = 0;
        int ckc = 0;                                             // columns since the last checkpoint
        Raw4 nx0 = load4(seq0, 0, lim0), nx1 = load4(seq1, 0, lim1);

        // one column; `general` (compile-time after inlining): candidate masks and final-column handling
        auto column = [&](const int j, const uint32_t c8_0, const uint32_t c8_1, const bool general) {
            const uint32_t *tp = tbl + pair_index(c8_0, c8_1);
            const uint32_t rowTp = rowT;
            rowT += Bp;
            uint32_t Ml;
            col_score<NR, EXACT>(Mo, Hl, tp, rowTp, rowT, GEB, Ml);
            uint32_t cand = Ml + candK - rowT;                    // scout, row class: strict '>' in visiting order (dp_scout.h:167-180)
            if (general)
                cand &= ((j >= lo0 && j <= hi0) ? 0x0000FFFFu : 0u) | ((j >= lo1 && j <= hi1) ? 0xFFFF0000u : 0u);
            bool keep_hi, keep_lo;
            bestS = vmax_keep(bestS, cand, keep_hi, keep_lo);
            if (!keep_lo) ord0 = j;
            if (!keep_hi) ord1 = j;
            if (general) {
                const bool d0 = fin0 && j == len0, d1 = fin1 && j == len1;
                if (d0 || d1) {
                    // final column of a half (adapter overhanging the read end): rows 1..n in order; V rebuilt as the
                    // forward pass computed it.  Through shared memory, so that the row loop is not unrolled.
#pragma unroll
                    for (int r = 0; r < NR; ++r) { s_fl[(2 * r) * PCB_BLOCK] = Mo[r]; s_fl[(2 * r + 1) * PCB_BLOCK] = Hl[r]; }
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (!(h ? d1 : d0)) continue;
                        const int forced = tk[h].forced, len = tk[h].len;
                        int Tn = half16(rowT, h) + b * r0, Vp = 0, Mp = Tn, best = 0, bi = 0, be = 0;     // the row above adapter row 1
                        for (int i = 1; i <= n; ++i) {
                            const int M = half16(s_fl[(2 * (r0 + i - 1)) * PCB_BLOCK], h), H = half16(s_fl[(2 * (r0 + i - 1) + 1) * PCB_BLOCK], h);
                            const int V = max(Vp + geb, Mp);
                            Tn += b;
                            const int c = M + OFF - Tn;
                            if ((forced < 0 || forced == len + i) && c > best) { best = c; bi = i; be = ((V == M) ? 2 : 0) | ((H == M) ? 1 : 0); }
                            Vp = V; Mp = M;
                        }
                        if (h) { finS1 = best; finI1 = bi; finE1 = be; } else { finS0 = best; finI0 = bi; finE0 = be; }
                    }
                }
            }
        };

        for (int j0 = 1; j0 <= mw; j0 += 4) {
            uint32_t cur0 = join4(nx0), cur1 = join4(nx1);
            nx0 = load4(seq0, j0 + 3, lim0); nx1 = load4(seq1, j0 + 3, lim1);       // used one group later
            if (TRACE && ckc == 0 && j0 > 1) {                   // checkpoint k = state after column j0 - 1 = k * C  (column 0 is known)
                // One 16-bit word per cell and task: M in 13 bits, M - H in 3.  H only matters to the next column while
                // H + ge >= M + go, i.e. M - H <= ge - go =: t; any H further down gives the same value and the same flag, so
                // M - H is stored clamped to t + 1 <= 7 (pcf: fast schemes have go >= -7, which also keeps M below 2^13).
                // (address from base + running offset: a pointer bumped after the stores waits for all of them to read it;
                //  streaming stores: a record is read back at most once, by another SM, much later.  Every store of the warp
                //  is one contiguous KB: records kept per thread -- 32 scattered 16-byte pieces per instruction -- held the
                //  shared-memory loads of the column loop up behind them in the memory pipe)
                uint32_t *const cq = ckp + ckoff;
                PCF_CHECK(L, (uint64_t)ckoff + ((NR + 7) / 8) * 256u <= 32u * ckw && unit - L.unit0 < L.unit1 - L.unit0);
                const uint32_t T1 = GEB + 0x00010001u;
#pragma unroll
                for (int g = 0; g < (NR + 7) / 8; ++g) {
                    uint32_t v[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) { const int r = 8 * g + e; v[e] = r < NR ? Mo[r < NR ? r : 0] + __vminu2(Mo[r < NR ? r : 0] - Hl[r < NR ? r : 0], T1) * 8192u : 0u; }
                    st256(cq + g * 256, v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
                }
                ckoff += ((NR + 7) / 8) * 256;
            }
            if (TRACE) { ckc += 4; if (ckc == C) ckc = 0; }
            if (!TRACE && ((j0 - 1) & (REBASE - 1)) == 0 && j0 > 1) {   // keep the growing bias inside 15 bits
                const uint32_t d = pk2(b * REBASE);
#pragma unroll
                for (int r = 0; r < NR; ++r) { Mo[r] -= d; Hl[r] -= d; }
                rowT -= d;
            }
            if (j0 >= ulo && j0 + 3 <= uhi) {
#pragma unroll UNR
                for (int k = 0; k < 4; ++k) { column(j0 + k, cur0 & 0xFFu, cur1 & 0xFFu, false); cur0 >>= 8; cur1 >>= 8; }
            } else {
#pragma unroll 1
                for (int k = 0; k < 4; ++k) { column(j0 + k, cur0 & 0xFFu, cur1 & 0xFFu, true); cur0 >>= 8; cur1 >>= 8; }
            }
        }

        // merge row class and final-column class (final-column cells are visited last: they win only if strictly better)
        int sc0 = (int)(bestS & 0xFFFFu), sc1 = (int)(bestS >> 16);
        bool fc0 = false, fc1 = false;
        if (finS0 > sc0) { sc0 = finS0; ord0 = len0 + finI0; fc0 = true; }
        if (finS1 > sc1) { sc1 = finS1; ord1 = len1 + finI1; fc1 = true; }

        if (!TRACE) {
            // order key as pcb_kernels.cu: (score << 16) | ((0x3FFF - ord) << 2); the eq bits are re-derived by the trace launch
            if (tk[0].out_idx >= 0) L.scan_out[tk[0].out_idx] = (len0 > 0 && sc0 > 0) ? (int)(((uint32_t)(sc0 - OFF) << 16) | ((uint32_t)(0x3FFF - ord0) << 2)) : (int)0x80000000;
            if (tk[1].out_idx >= 0) L.scan_out[tk[1].out_idx] = (len1 > 0 && sc1 > 0) ? (int)(((uint32_t)(sc1 - OFF) << 16) | ((uint32_t)(0x3FFF - ord1) << 2)) : (int)0x80000000;
            continue;
        }
        // every task that has an end cell becomes an entry for k_rounds
        uint32_t ex[2], ey[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const HalfTask &t = tk[h];
            const int sc = h ? sc1 : sc0, ord = h ? ord1 : ord0;
            const bool fc = h ? fc1 : fc0;
            ex[h] = 0u; ey[h] = 0u;
            if (t.out_idx < 0) continue;
            if (t.len <= 0) { write_failed(L.out, t.out_idx); continue; }
            if (sc <= 0) { atomicOr(L.err_flag, 2); write_failed(L.out, t.out_idx); continue; }   // forced cell not found
            int pci, pcj, pst;
            if (fc) { pci = ord - t.len; pcj = t.len; const int e = h ? finE1 : finE0; pst = (e & 2) ? 1 : ((e & 1) ? 2 : 0); }
            else    { pci = n; pcj = ord; pst = 4; }      // 4: the end cell's V == M / H == M still decide the start state
            ex[h] = (uint32_t)pci | ((uint32_t)pcj << 7) | ((uint32_t)pst << 17) | 0x80000000u;
            ey[h] = (uint32_t)sc;
        }
        *entp = make_uint4(ex[0], ey[0], ex[1], ey[1]);
    }
}

// ------------------------------------------------------------------------------------------------
// Pass 2 of the trace: sort, then rounds.
//
// A traceback crosses 1.8 flag blocks (C columns) on average; its first block is 1..C columns wide (7 on average: the
// end cell sits anywhere in it) and needs all adapter rows, every later block is C columns wide but only needs the rows
// above the row in which the path enters it (13 of 24 on average, 5 for a third block).  A warp recomputes a block for 64
// tracebacks in lockstep, so what a round costs is the LARGEST width x the LARGEST row count among them.  Hence:
//   generation 0 = the first block of every traceback, generation g = the next block of what is still under way, the last
//   generation (L.finish) = all blocks that are left; 2 to 5 generations (pcf_generations).  Before a generation runs, its
//   items are sorted (counting sort: k_count0 / in-kernel counts, k_bucket_scan, k_scatter) by adapter and then by
//   width (first blocks of paths that end in the last adapter row) or by rows (everything else), and a round takes 64
//   neighbours of that order: 300 instead of 690 flag cells per traceback.
//   A round walks each path to the left edge of its block; paths that end there are finished (statistics, result record),
//   the others go to the list of the next generation with the state of their walk.  A round of the last generation keeps
//   its 64 items until all of them are finished (sorted by rows, the ones that go on are neighbours).
// dynamic smem of k_rounds: [W][NR*25] score tables, [W][NR] adapter codes, then per thread (strided by PCB_BLOCK) C*NR/4
//   flag words.
// ------------------------------------------------------------------------------------------------
constexpr int NB = PCF_NB;
struct Meta {                      // views into L.meta (pcf_meta_words)
    int32_t *hist0, *hist1, *base, *cursor, *chunk_base, *counters;
};
__device__ __forceinline__ Meta meta_of(const PcbLaunch &L)
{
    Meta m; const int64_t A = L.n_major;
    m.hist0 = L.meta; m.hist1 = m.hist0 + A * NB; m.base = m.hist1 + A * NB; m.cursor = m.base + A * NB + 1;
    m.chunk_base = m.cursor + A * NB; m.counters = m.chunk_base + A + 1;
    return m;
}
// counters: 0 / 4 = items in `surv` after an even / odd generation, 1..3 and 5.. = work-queue heads of generations 0..2 and 3..
__device__ __forceinline__ int32_t major_of(const PcbLaunch &L, int64_t u /* unit - unit0 */, int64_t groups)
{
    return L.src.units ? L.src.units[L.unit0 + u].adapter : (int32_t)((L.unit0 + u) / groups);
}
// bucket of a first block: 0 = no block at all (empty path), 1..C = its width for paths that end in the last adapter row,
// 17 + (rows - 1) / 4 for paths that end in the read's final column (fewer rows, all of one width)
// (skew = 1 for the two-stripe kernels, whose last adapter row runs one iteration behind its column)
__device__ __forceinline__ int bucket0(uint32_t x, int C, int skew)
{
    const int ci = (int)(x & 127u), cj = (int)((x >> 7) & 0x3FFu), st = (int)((x >> 17) & 7u);
    if (ci == 0 || cj == 0) return 0;
    return st == 4 ? cj + skew - ((cj + skew - 1) / C) * C : min(NB - 1, 17 + ((ci - 1) >> 2));
}
__device__ __forceinline__ int bucket1(uint32_t pos) { return (int)(((pos & 127u) - 1u) >> 2); }
// two-stripe kernels: a path is over when it reaches row 0 or column 0, and with > 72 rows the columns often run out first;
// rounds of equal remaining length finish together
__device__ __forceinline__ int bucket1_two(uint32_t pos) { return (int)((min(pos & 127u, (pos >> 7) & 0x3FFu) - 1u) >> 2); }

// one thread per (unit, lane): counts the first blocks per (major, bucket)
__global__ void k_count0(const PcbLaunch L, int C, int skew)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nu = L.unit1 - L.unit0;
    if ((idx >> 5) >= nu) return;                                     // whole warps leave together
    const Meta M = meta_of(L);
    const int64_t groups = L.src.units ? 0 : (L.src.n_windows + 63) / 64;
    const int lane = threadIdx.x & 31;
    const uint4 e = L.entries[idx];
    const int32_t a = major_of(L, idx >> 5, groups);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t x = h ? e.z : e.x;
        const int key = (x >> 31) ? bucket0(x, C, skew) : -1;
        const uint32_t m = __match_any_sync(0xffffffffu, key);
        if (key >= 0 && lane == __ffs(m) - 1) atomicAdd(M.hist0 + (int64_t)a * NB + key, __popc(m));
    }
}

// one CTA: bucket counts -> first slot of every (major, bucket) in the sorted order, and the chunk table of the rounds
// (a chunk = 64 consecutive items of one major).  Clears the counts for their next use.
__global__ void k_bucket_scan(const PcbLaunch L, int which)
{
    const Meta M = meta_of(L);
    int32_t *const hist = which ? M.hist1 : M.hist0;
    const int A = L.n_major, tid = threadIdx.x;
    __shared__ int s_n[1024], s_c[1024];
    __shared__ int carry_n, carry_c;
    if (tid == 0) { carry_n = 0; carry_c = 0; }
    __syncthreads();
    for (int a0 = 0; a0 < A; a0 += 1024) {
        const int a = a0 + tid;
        int n = 0;
        if (a < A) for (int b = 0; b < NB; ++b) n += hist[(int64_t)a * NB + b];
        s_n[tid] = n; s_c[tid] = (n + 63) >> 6;
        __syncthreads();
        for (int d = 1; d < 1024; d <<= 1) {                          // inclusive scans
            const int vn = tid >= d ? s_n[tid - d] : 0, vc = tid >= d ? s_c[tid - d] : 0;
            __syncthreads();
            s_n[tid] += vn; s_c[tid] += vc;
            __syncthreads();
        }
        if (a < A) {
            int run = carry_n + s_n[tid] - n;
            M.chunk_base[a] = carry_c + s_c[tid] - ((n + 63) >> 6);
            for (int b = 0; b < NB; ++b) {
                const int64_t k = (int64_t)a * NB + b;
                M.base[k] = run; M.cursor[k] = 0; run += hist[k]; hist[k] = 0;
            }
        }
        __syncthreads();
        if (tid == 1023) { carry_n += s_n[1023]; carry_c += s_c[1023]; }
        __syncthreads();
    }
    // (the survivor counter the NEXT rounds launch fills: 0 after a scan of hist0, 4 after a scan of hist1)
    if (tid == 0) { M.base[(int64_t)A * NB] = carry_n; M.chunk_base[A] = carry_c; M.counters[which == 0 ? 0 : 4] = 0; }
}

__device__ __forceinline__ void put_item(PcfItem *dst, uint32_t ref, uint32_t pos, uint32_t endc, uint32_t a0, uint32_t a1, uint32_t a2, uint64_t off)
{
    st256(reinterpret_cast<uint32_t *>(dst), ref, pos, endc, a0, a1, a2, (uint32_t)off, (uint32_t)(off >> 32));
}
// slot of an item of `key` in the sorted order (one atomic per distinct key of the warp)
__device__ __forceinline__ int64_t sorted_slot(const Meta &M, int key, int lane)
{
    const uint32_t m = __match_any_sync(0xffffffffu, key);
    const int leader = __ffs(m) - 1;
    int first = 0;
    if (key >= 0 && lane == leader) first = atomicAdd(M.cursor + key, __popc(m));
    first = __shfl_sync(0xffffffffu, first, leader);
    return key >= 0 ? (int64_t)M.base[key] + first + __popc(m & ((1u << lane) - 1u)) : -1;
}
// generation 0: one thread per (unit, lane); the end cells of pass 1 become items, in sorted order
__global__ void k_scatter0(const PcbLaunch L, int C, int skew)
{
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nu = L.unit1 - L.unit0;
    if ((idx >> 5) >= nu) return;
    const Meta M = meta_of(L);
    const int64_t groups = L.src.units ? 0 : (L.src.n_windows + 63) / 64;
    const int lane = threadIdx.x & 31;
    const uint4 e = L.entries[idx];
    const int32_t a = major_of(L, idx >> 5, groups);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const uint32_t x = h ? e.z : e.x, y = h ? e.w : e.y;
        const int key = (x >> 31) ? a * NB + bucket0(x, C, skew) : -1;
        const int64_t slot = sorted_slot(M, key, lane);
        PCF_CHECK(L, slot < 64 * nu);
        if (slot >= 0) {
            // where the task's bases are: pass 2 then starts from the item alone
            const int64_t unit = L.unit0 + (idx >> 5);
            const uint64_t off = L.src.units ? L.src.tasks[L.src.units[unit].task0 + 2 * lane + h].off
                                             : (uint64_t)((unit - (int64_t)a * groups) * 64 + 2 * lane + h) * (uint64_t)L.src.stride;
            put_item(L.sorted + slot, ((uint32_t)(idx >> 5) << 6) | (uint32_t)(2 * lane + h), x & 0xFFFFFu, (x & 0x1FFFFu) | (y << 17), 0u, 0u, 0u, off);
        }
    }
}
// later generations: one thread per surviving item (`counter`: which of the two survivor counts holds their number)
__global__ void k_scatter1(const PcbLaunch L, int two, int counter)
{
    const Meta M = meta_of(L);
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, n = M.counters[counter];
    if ((idx & ~(int64_t)31) >= n) return;
    const int64_t groups = L.src.units ? 0 : (L.src.n_windows + 63) / 64;
    const int lane = threadIdx.x & 31;
    uint32_t v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int key = -1;
    if (idx < n) {
        ld256(reinterpret_cast<const uint32_t *>(L.surv + idx), v);
        key = major_of(L, v[0] >> 6, groups) * NB + (two ? bucket1_two(v[1]) : bucket1(v[1]));
    }
    const int64_t slot = sorted_slot(M, key, lane);
    PCF_CHECK(L, slot < 64 * (L.unit1 - L.unit0) && n <= 64 * (L.unit1 - L.unit0));
    if (slot >= 0) put_item(L.sorted + slot, v[0], v[1], v[2], v[3], v[4], v[5], (uint64_t)v[6] | ((uint64_t)v[7] << 32));
}

// The walk's state for one traceback (dp_traceback_impl.h:379-451): i, j = cell the walk stands in; s = state, 0 = M, 1 = V,
// 2 = H; the accumulators (align_stats.h: WalkAcc) are packed so that a step is branch-free.
struct Walk { int i, j, s; WalkAcc acc; };
__device__ __forceinline__ void walk_unpack(Walk &w, uint32_t pos, uint32_t a0, uint32_t a1, uint32_t a2)
{
    w.i = (int)(pos & 127u); w.j = (int)((pos >> 7) & 0x3FFu); w.s = (int)((pos >> 17) & 7u); w.acc.op = (int)(pos >> 20);
    w.acc.kAR = a0; w.acc.tHV = a1; w.acc.nHV = a2;
}
__device__ __forceinline__ uint32_t walk_pos(const Walk &w)
{
    return (uint32_t)w.i | ((uint32_t)w.j << 7) | ((uint32_t)w.s << 17) | ((uint32_t)w.acc.op << 20);
}
// One op of the reverse walk; flag word layout as col_flags writes it.  ar = register row of the walk's adapter row, jc = its
// column inside the flag block, pos2 = (2 * column + 1) | (2 * row + 1) << 16.  In state M a cell whose diagonal flag is off
// hands over to V or H and that state's step is taken at once, from the same cell's flags, as SeqAn does.  No branches:
// the two tracebacks of a thread are stepped together and their shared-memory loads overlap.
template <int FW>
__device__ __forceinline__ uint32_t walk_flags(const uint32_t *s_fl, const int h, const int ar, const int jc)
{
    return s_fl[(jc * FW + (ar >> 2)) * PCB_BLOCK] >> (8 * h + 2 * (ar & 3));
}
template <int FW>
__device__ __forceinline__ void walk_step(const uint32_t *s_fl, const int h, int &ar, int &jc, uint32_t &pos2, Walk &w)
{
    const uint32_t f = walk_flags<FW>(s_fl, h, ar, jc);
#ifdef PCF_CHECKED
    if (ar < 0 || (ar >> 2) >= FW || jc < 0 || jc >= 16) asm volatile("trap;");
#endif
    const uint32_t fh = f & 1u, fg = (f >> 1) & 1u, fv = (f >> 16) & 1u, fd = (f >> 17) & 1u;
    const uint32_t sM = w.s == 0 ? 1u : 0u, sV = w.s == 1 ? 1u : 0u;
    const uint32_t isD = sM & fd, isV = sV | (sM & (fd ^ 1u) & fg), isH = (isD | isV) ^ 1u;
    const uint32_t extV = isV & fv, extH = isH & fh;
    walk_acc_step(w.acc, pos2, isD, isV, isH, extV | extH);
    w.s = (int)(extV | (extH << 1));
    ar -= (int)(isH ^ 1u); jc -= (int)(isV ^ 1u);
}

template <int NR, bool EXACT, int C>
__global__ void __launch_bounds__(PCB_BLOCK, PCF_MINB_P2(NR)) k_rounds(const PcbLaunch L)
{
    extern __shared__ uint32_t smem[];
    constexpr int FW = NR / 4, W = PCB_BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *const tbl = smem + warp * (NR * TP);
    uint32_t *const s_code = smem + W * (NR * TP) + warp * NR;
    uint32_t *const s_fl = smem + W * (NR * TP + NR) + threadIdx.x;

    const Meta M = meta_of(L);
    const int A = L.n_major;
    const bool finish = L.finish != 0;                          // the last generation: a round keeps its items until all are done
    const int b = -L.sc.gap_open;
    const int s_match = L.sc.match + 2 * b, s_mis = L.sc.mismatch + 2 * b;
    const uint32_t GEB = pk2(L.sc.gap_extend + b), Bp = pk2(b);
    const bool cross = L.src.units == nullptr;
    const int64_t groups = cross ? (L.src.n_windows + 63) / 64 : 0;
    const uint64_t ckw = ckw_of(L.mmax, C, NR);                  // checkpoint words per thread and unit

    // Which chunks a warp takes.  NR = 24 (the class that is cut into concurrent pipelines, whose kernels fill each other's
    // idle SMs): an equal share of the chunk list.  Larger NR: chunks are handed out from a queue head in meta, the LAST
    // chunks first -- inside a major the sort puts the most expensive rounds last, so they start early and the cheap ones
    // fill the tail of the launch (measured: 25..32 bases 0.89 -> 0.81 ms, 33..40 bases 0.85 -> 0.74 ms; 24 bases 13.2 ->
    // 13.7 ms, hence the split).
#ifdef PCF_DYN_ALL
    constexpr bool DYN = true;
#else
    constexpr bool DYN = NR > 24;
#endif
    const int64_t nwarps = (int64_t)gridDim.x * W, gw = (int64_t)blockIdx.x * W + warp;
    const int64_t total = M.chunk_base[A];
    int64_t c_cur = DYN ? 0 : gw * total / nwarps;
    const int64_t c_end = DYN ? 0 : (gw + 1) * total / nwarps;
    int32_t *const qhead = M.counters + (L.gen < 3 ? 1 : 2) + L.gen;
    int a = -1, n = 0, r0 = 0;
    int32_t tbl_adapter = -1, out_col = 0;
    int64_t a_item0 = 0, a_item1 = 0, a_chunk0 = 0, a_chunk1 = 0;

    for (;;) {
        int64_t c;
        if (DYN) {
            int64_t q = 0;
            if (lane == 0) q = atomicAdd(qhead, 1);
            q = __shfl_sync(0xffffffffu, q, 0);
            if (q >= total) break;
            c = total - 1 - q;
        } else {
            if (c_cur >= c_end) break;
            c = c_cur++;
        }
        if (a < 0 || c < a_chunk0 || c >= a_chunk1) {             // another major: the one whose chunk range holds c
            int lo = 0, hi = A - 1;
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (M.chunk_base[mid] <= c) lo = mid; else hi = mid - 1; }
            a = lo; a_chunk1 = M.chunk_base[a + 1];
            a_item0 = M.base[(int64_t)a * NB]; a_item1 = M.base[(int64_t)(a + 1) * NB]; a_chunk0 = M.chunk_base[a];
            const int32_t adapter = cross ? L.src.ids[a] : a;
            if (cross) out_col = L.src.id_pos ? L.src.id_pos[a] : a;
            const int64_t aoff = L.ad.off[adapter];
            n = (int)(L.ad.off[adapter + 1] - aoff);
            r0 = EXACT ? 0 : NR - n;
            if (adapter != tbl_adapter) {
                __syncwarp();
                for (int i = lane; i < NR; i += 32) s_code[i] = i >= r0 ? (uint32_t)(L.ad.codes[aoff + i - r0] / PCB_CODE_SCALE) : 0u;
                __syncwarp();
                for (int e = lane; e < NR * TP; e += 32) {
                    const int r = e / TP, p = e - r * TP;
                    const uint32_t ac = s_code[r], c0 = (uint32_t)(p / 5), c1 = (uint32_t)(p % 5);
                    tbl[e] = r >= r0 ? pk(c0 == ac ? s_match : s_mis, c1 == ac ? s_match : s_mis) : pk(2 * b, 2 * b);
                }
                __syncwarp();
                tbl_adapter = adapter;
            }
        }
        // this lane's two items
        const int64_t i0 = a_item0 + (c - a_chunk0) * 64 + 2 * lane;
        bool valid[2];
        uint32_t ref[2], endc[2];
        Walk wk[2];
        uint64_t off[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            valid[h] = i0 + h < a_item1;
            PCF_CHECK(L, a_item1 <= 64 * (L.unit1 - L.unit0) && a_item0 <= a_item1);
            uint32_t v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (valid[h]) ld256(reinterpret_cast<const uint32_t *>(L.sorted + i0 + h), v);
            ref[h] = v[0]; endc[h] = v[2];
            walk_unpack(wk[h], v[1], v[3], v[4], v[5]);
            off[h] = (uint64_t)v[6] | ((uint64_t)v[7] << 32);
        }
        // (bases are fetched a few bytes past a segment's end: the arenas are padded, and columns past the end never count)
        const uint8_t *const seq0 = L.arena + off[0], *const seq1 = L.arena + off[1];
        constexpr int lim0 = 0x7fffffff, lim1 = 0x7fffffff;

        for (;;) {
            // block of each half: columns k*C+1 .. k*C+C, holding its current column
            const bool go0 = valid[0] && wk[0].i > 0 && wk[0].j > 0, go1 = valid[1] && wk[1].i > 0 && wk[1].j > 0;
            const int k0 = go0 ? (wk[0].j - 1) / C : 0, k1 = go1 ? (wk[1].j - 1) / C : 0;
            const int jb0 = k0 * C, jb1 = k1 * C;
            const int tmax = __reduce_max_sync(0xffffffffu, max(go0 ? wk[0].j - jb0 : 0, go1 ? wk[1].j - jb1 : 0));
            // register rows the flags are needed for: everything above the rows the walks stand in
            const int rlim = __reduce_max_sync(0xffffffffu, max(go0 ? (r0 + wk[0].i + 3) & ~3 : 0, go1 ? (r0 + wk[1].i + 3) & ~3 : 0));
            if (tmax > 0) {
                uint32_t q0[C / 4], q1[C / 4];                   // the block's read bases
#pragma unroll
                for (int g = 0; g < C / 4; ++g) { q0[g] = join4(load4(seq0, jb0 + 4 * g, lim0)); q1[g] = join4(load4(seq1, jb1 + 4 * g, lim1)); }
                uint32_t Mo[NR], Hl[NR];
                {
                    // the checkpointed column the block starts from: written by the thread that ran the task's pass 1
                    // (slot >> 1), two tasks per word, M | (M - H) << 13 per task; column 0 is not stored (M = 0, H = -inf)
                    constexpr int G8 = (NR + 7) / 8;
                    const uint32_t *p0 = L.scratch + ((uint64_t)(ref[0] >> 6) * 32u * ckw + (uint64_t)max(k0 - 1, 0) * G8 * 256u + (uint64_t)((ref[0] & 63u) >> 1) * 8u);
                    const uint32_t *p1 = L.scratch + ((uint64_t)(ref[1] >> 6) * 32u * ckw + (uint64_t)max(k1 - 1, 0) * G8 * 256u + (uint64_t)((ref[1] & 63u) >> 1) * 8u);
                    const uint32_t sel = ((ref[0] & 1u) ? 0x0032u : 0x0010u) | ((ref[1] & 1u) ? 0x7600u : 0x5400u);
                    const uint32_t T1s = (GEB + 0x00010001u) * 8192u;
                    const bool ld0 = go0 && k0 > 0, ld1 = go1 && k1 > 0;
                    PCF_CHECK(L, !ld0 || ((int64_t)(ref[0] >> 6) < L.unit1 - L.unit0 && (uint64_t)k0 * G8 * 8u <= ckw));
                    PCF_CHECK(L, !ld1 || ((int64_t)(ref[1] >> 6) < L.unit1 - L.unit0 && (uint64_t)k1 * G8 * 8u <= ckw));
                    PCF_CHECK(L, tmax <= C && rlim <= NR);
#pragma unroll
                    for (int g = 0; g < G8; ++g) {
                        if (8 * g < rlim) {
                            uint32_t x[8], y[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) x[e] = y[e] = pk2(B0 + b * (8 * g + e + 1)) + T1s;
                            if (ld0) ld256(p0 + g * 256, x);
                            if (ld1) ld256(p1 + g * 256, y);
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                if (8 * g + e < NR) {
                                    const uint32_t v = prmt(x[e], y[e], sel), m = v & 0x1FFF1FFFu;
                                    Mo[8 * g + e] = m; Hl[8 * g + e] = m - ((v >> 13) & 0x00070007u);
                                }
                            }
                        } else {
#pragma unroll
                            for (int e = 0; e < 8; ++e) if (8 * g + e < NR) { Mo[8 * g + e] = 0u; Hl[8 * g + e] = 0u; }
                        }
                    }
                }
                uint32_t rT = pk(B0 + b * jb0, B0 + b * jb1);
                const bool pend0 = go0 && (wk[0].s & 4), pend1 = go1 && (wk[1].s & 4);
                const bool anypend = __any_sync(0xffffffffu, pend0 || pend1);
                int s0 = wk[0].s & 3, s1 = wk[1].s & 3;
#pragma unroll UNR2
                for (int t = 1; t <= tmax; ++t) {
                    const uint32_t *tp = tbl + pair_index(q0[0] & 0xFFu, q1[0] & 0xFFu);
#pragma unroll
                    for (int g = 0; g < C / 4; ++g) {                  // shift the base queues by one byte
                        q0[g] = __funnelshift_r(q0[g], g + 1 < C / 4 ? q0[(g + 1) % (C / 4)] : 0u, 8);
                        q1[g] = __funnelshift_r(q1[g], g + 1 < C / 4 ? q1[(g + 1) % (C / 4)] : 0u, 8);
                    }
                    const uint32_t rowTp = rT;
                    rT += Bp;
                    uint32_t Ml, Vl, Hlast;
                    col_flags<NR, EXACT>(Mo, Hl, tp, rowTp, rT, GEB, s_fl + (t - 1) * FW * PCB_BLOCK, rlim, Ml, Vl, Hlast);
                    if (anypend) {
                        // end cell in the last adapter row: V == M / H == M of that cell decide how the trace starts
                        // (_correctTraceValue, dp_algorithm_impl.h:1169-1185)
                        const uint32_t xv = Ml ^ Vl, xh = Ml ^ Hlast;
                        if (pend0 && jb0 + t == wk[0].j) s0 = ((xv & 0xFFFFu) == 0) ? 1 : (((xh & 0xFFFFu) == 0) ? 2 : 0);
                        if (pend1 && jb1 + t == wk[1].j) s1 = ((xv >> 16) == 0) ? 1 : (((xh >> 16) == 0) ? 2 : 0);
                    }
                }
                __syncwarp();
                wk[0].s = s0; wk[1].s = s1;                              // the pending bit is spent
                // walk (dp_traceback_impl.h:379-451, :498-545) inside the block: both halves together while both are on their way
                {
                    Walk w0 = wk[0], w1 = wk[1];
                    int ar0 = r0 + w0.i - 1, jc0 = w0.j - jb0 - 1, ar1 = r0 + w1.i - 1, jc1 = w1.j - jb1 - 1;
                    uint32_t pz0 = (uint32_t)(2 * w0.j + 1) | ((uint32_t)(2 * w0.i + 1) << 16), pz1 = (uint32_t)(2 * w1.j + 1) | ((uint32_t)(2 * w1.i + 1) << 16);
                    bool act0 = go0, act1 = go1;
                    while (act0 && act1) {
                        walk_step<FW>(s_fl, 0, ar0, jc0, pz0, w0); walk_step<FW>(s_fl, 1, ar1, jc1, pz1, w1);
                        act0 = ar0 >= r0 && jc0 >= 0; act1 = ar1 >= r0 && jc1 >= 0;
                    }
                    while (act0) { walk_step<FW>(s_fl, 0, ar0, jc0, pz0, w0); act0 = ar0 >= r0 && jc0 >= 0; }
                    while (act1) { walk_step<FW>(s_fl, 1, ar1, jc1, pz1, w1); act1 = ar1 >= r0 && jc1 >= 0; }
                    if (go0) { w0.i = ar0 - r0 + 1; w0.j = jc0 + jb0 + 1; }
                    if (go1) { w1.i = ar1 - r0 + 1; w1.j = jc1 + jb1 + 1; }
                    wk[0] = w0; wk[1] = w1;
                }
                __syncwarp();                                            // the flag words are rewritten by the next round
            }

            // paths that reached row 0 / column 0 are complete (this also covers the empty path)
            bool keep[2];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const Walk &w = wk[h];
                keep[h] = valid[h] && w.i > 0 && w.j > 0;
                if (!valid[h] || keep[h]) continue;
                valid[h] = false;
                int col0 = 0, mtrue;
                int64_t oidx;
                if (cross) {
                    const int64_t wi = (L.unit0 + (int64_t)(ref[h] >> 6) - (int64_t)a * groups) * 64 + (int64_t)(ref[h] & 63u);
                    mtrue = L.src.win_len[wi]; oidx = wi * L.src.n_ids_total + out_col;
                } else {
                    const PcbTask &k = L.src.tasks[L.src.units[L.unit0 + (int64_t)(ref[h] >> 6)].task0 + (int64_t)(ref[h] & 63u)];
                    col0 = k.col0; mtrue = k.m_true; oidx = k.out_idx;
                }
                const int endi = (int)(endc[h] & 127u), endj = (int)((endc[h] >> 7) & 0x3FFu), score = (int)(endc[h] >> 17) - OFF;
                if (w.j == 0 && w.i > 0 && col0 > 0) atomicOr(L.err_flag, 1);      // window too narrow for this path
                // the path as align_stats.h wants it.  Matches are not counted on the way: the path's ops and the end
                // cell's score determine them (every D op scores match or mismatch, every gap of g ops go + (g-1) ge)
                int ok;
                const int matches = walk_acc_matches(w.acc, endi, endj, w.i, w.j, score, L.sc.match, L.sc.mismatch, L.sc.gap_open, L.sc.gap_extend, &ok);
                if (!ok) atomicOr(L.err_flag, 16);
                PathAcc pa;
                walk_acc_finish(w.acc, endi, endj, w.i, w.j, matches, pa);
                PathStats ps;
                path_finalize(pa, (int64_t)col0 + w.j, w.i, (int64_t)mtrue - (col0 + endj), n - endi, mtrue, n, ps);
                pcb_result res;
                res.read_start = ps.read_start; res.read_end = ps.read_end;
                res.adapter_start = ps.adapter_start; res.adapter_end = ps.adapter_end;
                res.raw_score = score; res.matches = ps.matches;
                res.aligned_len = ps.aligned_len; res.full_len = ps.full_len;
                PCF_CHECK(L, oidx >= 0 && (L.n_out <= 0 || oidx < L.n_out));
                L.out[oidx] = res;
            }
            if (finish) {
                if (!__any_sync(0xffffffffu, keep[0] || keep[1])) break;
                continue;                                                // these 64 stay together until the last is done
            }
            // the others go to the next generation, counted per (major, rows): even generations count in hist1 / counters[0],
            // odd ones in hist0 / counters[4] (free again by then: k_bucket_scan clears what it has read)
            int32_t *const n_surv = M.counters + ((L.gen & 1) ? 4 : 0), *const hist_next = (L.gen & 1) ? M.hist0 : M.hist1;
            const uint32_t m0 = __ballot_sync(0xffffffffu, keep[0]), m1 = __ballot_sync(0xffffffffu, keep[1]);
            const int nk = __popc(m0) + __popc(m1);
            if (nk) {
                int first = 0;
                if (lane == 0) first = atomicAdd(n_surv, nk);
                first = __shfl_sync(0xffffffffu, first, 0);
                const uint32_t lt = (1u << lane) - 1u;
                PCF_CHECK(L, first + nk <= 64 * (L.unit1 - L.unit0));
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const Walk &w = wk[h];
                    const uint32_t pos = walk_pos(w);
                    const int key = keep[h] ? a * NB + bucket1(pos) : -1;
                    const uint32_t mk = __match_any_sync(0xffffffffu, key);
                    if (keep[h]) {
                        if (lane == __ffs(mk) - 1) atomicAdd(hist_next + key, __popc(mk));
                        const int slot = first + (h ? __popc(m0) + __popc(m1 & lt) : __popc(m0 & lt));
                        put_item(L.surv + slot, ref[h], pos, endc[h], w.acc.kAR, w.acc.tHV, w.acc.nHV, off[h]);
                    }
                }
            }
            break;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Two-stripe kernels for adapters of 73..127 bases (the rapid full-sequence barcode sets, adapters.py:479-498: 102 and 111).
//
// 144 adapter rows do not fit one thread's registers as two tasks x 72 rows -- but they do as ONE task x two stripes of 72
// rows: the low halves of the packed registers hold rows 1..72, the high halves rows 73..144, and the high half runs one
// column BEHIND the low half (iteration t: low half = column t, high half = column t - 1).  What row 73 needs from row 72
// -- M(72, t-1), M(72, t-2), V(72, t-1) -- are low-half values of the two iterations before, kept in three registers and
// moved into the high half with a PRMT.  Every DPX instruction still does two cells, the code per row is the one-stripe
// code, and the substitution table simply holds (score(base[t], adapter[r]), score(base[t-1], adapter[72 + r])).
// A lane's two tasks of a unit are run one after the other; flag words, checkpoints and the walk use "iteration"
// t = column + stripe wherever the one-stripe kernels use the column.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t swap16(uint32_t x) { return prmt(x, x, 0x1032u); }
__device__ __forceinline__ uint32_t lo_hi(uint32_t lo_src, uint32_t hi_src) { return prmt(lo_src, hi_src, 0x7610u); }   // low half of the first, high half of the second

// col_score with the row above the first register row given: Din = M diagonal of that row (biased), Vu0 = V of the first row
template <int NR>
__device__ __forceinline__ void col_score2(uint32_t (&Mo)[NR], uint32_t (&Hl)[NR], const uint32_t *tp,
                                           uint32_t Din, uint32_t Vu0, uint32_t GEB, uint32_t &Ml, uint32_t &Vl)
{
    uint32_t Dn = Din + tp[0];
    uint32_t Vu = Vu0;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const uint32_t Hn = __viaddmax_s16x2(Hl[r], GEB, Mo[r]);
        if (r > 0) Vu = __viaddmax_s16x2(Vu, GEB, Mo[r - 1]);
        const uint32_t D = Dn;
        if (r < NR - 1) Dn = Mo[r] + tp[(r + 1) * TP];
        Mo[r] = __vimax3_s16x2(D, Vu, Hn); Hl[r] = Hn;
    }
    Ml = Mo[NR - 1]; Vl = Vu;
}

// col_flags with the row above given: Din as above, Mup0 / Vup0 = M and V of the row above the first register row
template <int NR>
__device__ __forceinline__ void col_flags2(uint32_t (&Mo)[NR], uint32_t (&Hl)[NR], const uint32_t *tp,
                                           uint32_t Din, uint32_t Mup0, uint32_t Vup0, uint32_t GEB, uint32_t *fp, const int rlim,
                                           uint32_t &Ml, uint32_t &Vl, uint32_t &Hlast)
{
    constexpr int PF = PCF_PF;
    uint32_t tv[PF];
#pragma unroll
    for (int d = 0; d < PF; ++d) tv[d] = tp[d * TP];
    uint32_t Dn = Din + tv[0];
    uint32_t Vu = Vup0, Wacc = 0;
    const uint32_t GEBG = GEB + GUARD;
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        if ((r & 3) == 0 && r > 0 && r >= rlim) break;
        const uint32_t D = Dn;
        if (r + PF < NR) tv[r % PF] = tp[(r + PF) * TP];
        const uint32_t Mup = (r == 0) ? Mup0 : Mo[(r > 0) ? r - 1 : 0];
        const uint32_t ch = Hl[r] + GEBG - Mo[r], cv = Vu + GEBG - Mup;
        const uint32_t Hn = __viaddmax_s16x2(Hl[r], GEB, Mo[r]), Vn = __viaddmax_s16x2(Vu, GEB, Mup);
        const uint32_t G = __vmaxs2(Vn, Hn);
        const uint32_t W = (sign4(ch, cv) & 0x55555555u) | (sign4(cmpge(Vn, Hn), cmpge(D, G)) & 0xAAAAAAAAu);
        if ((r & 3) == 0) Wacc = W;
        else if ((r & 3) == 1) Wacc = (Wacc & 0x03030303u) | (W & 0xFCFCFCFCu);
        else if ((r & 3) == 2) Wacc = (Wacc & 0x0F0F0F0Fu) | (W & 0xF0F0F0F0u);
        else Wacc = (Wacc & 0x3F3F3F3Fu) | (W & 0xC0C0C0C0u);
        if ((r & 3) == 3 || r == NR - 1) fp[(r >> 2) * PCB_BLOCK] = Wacc;
        if (r < NR - 1) Dn = Mo[r] + tv[(r + 1) % PF];
        Vu = Vn; Mo[r] = __vmaxs2(D, G); Hl[r] = Hn;
    }
    Ml = Mo[NR - 1]; Vl = Vu; Hlast = Hl[NR - 1];
}

// checkpoint words per thread and unit of the two-stripe kernels: 2 tasks x checkpoints x (row chunks + 1 chunk of boundary registers)
__device__ __forceinline__ uint64_t ckw2_of(int mmax, int C, int NR) { return 2ull * (uint64_t)((mmax + 1 + C - 1) / C) * (uint64_t)((NR + 7) / 8 + 1) * 8ull; }

// Score table of the two stripes: tbl[r * 25 + 5 * code(base[t]) + code(base[t-1])] = (score vs adapter row r, score vs row NR + r)
template <int NR>
__device__ __forceinline__ void build_table2(const PcbLaunch &L, uint32_t *tbl, uint32_t *s_code, int64_t aoff, int n, int r0, int lane,
                                             int s_match, int s_mis, int b)
{
    __syncwarp();
    for (int i = lane; i < 2 * NR; i += 32) s_code[i] = i >= r0 ? (uint32_t)(L.ad.codes[aoff + i - r0] / PCB_CODE_SCALE) : 0u;
    __syncwarp();
    for (int e = lane; e < NR * TP; e += 32) {
        const int r = e / TP, p = e - r * TP;
        const uint32_t a_lo = s_code[r], a_hi = s_code[NR + r], c0 = (uint32_t)(p / 5), c1 = (uint32_t)(p % 5);
        const int lo = r >= r0 ? (c0 == a_lo ? s_match : s_mis) : 2 * b;        // rows above the adapter: copies of DP row 0
        tbl[e] = pk(lo, c1 == a_hi ? s_match : s_mis);
    }
    __syncwarp();
}

// k_fast2<MODE, NR, C>: the score-only pass of the two-stripe kernels (MODE 0: scan keys, MODE 1: pass 1 of the trace).
// Same launch interface, unit / task mapping, entries and scan keys as k_fast.  Checkpoint k = state after iteration
// k * C: [unit - unit0][task & 1][k - 1][row chunks of 8 | boundary registers][lane][8 words].
// dynamic smem: [W][NR*25] score tables, [W][2*NR] adapter codes, then 2*NR words per thread for the final-column dump.
template <int MODE, int NR, int C>
__global__ void __launch_bounds__(PCB_BLOCK, 2) k_fast2(const PcbLaunch L)
{
    constexpr bool TRACE = MODE != 0;
    constexpr int R2 = 2 * NR, G8 = (NR + 7) / 8;
    extern __shared__ uint32_t smem[];
    constexpr int W = PCB_BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *const tbl = smem + warp * (NR * TP);
    uint32_t *const s_code = smem + W * (NR * TP) + warp * R2;
    uint32_t *const s_fl = smem + W * (NR * TP + R2) + threadIdx.x;
    const uint64_t ckw = ckw2_of(L.mmax, C, NR);

    const int b = -L.sc.gap_open;
    const uint32_t GEB = pk2(L.sc.gap_extend + b), Bp = pk2(b);
    const int s_match = L.sc.match + 2 * b, s_mis = L.sc.mismatch + 2 * b;
    const int geb = L.sc.gap_extend + b;

    const int64_t nwarps = (int64_t)gridDim.x * W, gw = (int64_t)blockIdx.x * W + warp;
    const bool cross = L.src.units == nullptr;
    const int64_t groups = cross ? (L.src.n_windows + 63) / 64 : 0;
    const int64_t nu = L.unit1 - L.unit0;
    int64_t u_begin, u_end, u_step;
    if (cross) { u_begin = L.unit0 + gw * nu / nwarps; u_end = L.unit0 + (gw + 1) * nu / nwarps; u_step = 1; }
    else       { u_begin = L.unit0 + gw; u_end = L.unit1; u_step = nwarps; }
    int32_t tbl_adapter = -1;

    for (int64_t unit = u_begin; unit < u_end; unit += u_step) {
        int32_t adapter = 0;
        HalfTask tk[2];
        if (cross) {
            const int64_t k = unit / groups, g = unit - k * groups;
            adapter = L.src.ids[k];
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                HalfTask &t = tk[h];
                t.off = 0; t.len = 0; t.col0 = 0; t.m_true = 0; t.forced = -1; t.own0 = 0; t.out_idx = -1;
                const int64_t w = g * 64 + 2 * lane + h;
                if (w < L.src.n_windows) {
                    t.off = (uint64_t)w * (uint64_t)L.src.stride;
                    t.len = L.src.win_len[w]; t.m_true = t.len;
                    t.out_idx = w * L.src.n_ids_total + (L.src.id_pos ? L.src.id_pos[k] : (int32_t)k);
                }
            }
        } else {
            fetch_task(L.src, unit, 2 * lane, adapter, tk[0]);
            fetch_task(L.src, unit, 2 * lane + 1, adapter, tk[1]);
        }
        uint4 *const entp = TRACE ? L.entries + (unit - L.unit0) * 32 + lane : nullptr;
        const int64_t aoff = L.ad.off[adapter];
        const int n = (int)(L.ad.off[adapter + 1] - aoff);
        if (n <= NR || n > R2 || n > 127) {                              // the host never sends these here; fail loudly if it does
            atomicOr(L.err_flag, 4);
            for (int h = 0; h < 2; ++h) if (tk[h].out_idx >= 0) {
                if (TRACE) write_failed(L.out, tk[h].out_idx); else L.scan_out[tk[h].out_idx] = (int)0x80000000;
            }
            if (TRACE) *entp = make_uint4(0u, 0u, 0u, 0u);
            continue;
        }
        const int r0 = R2 - n;                                 // bottom-aligned: adapter row i = register row r0 + i - 1 of 0 .. 2*NR - 1
        if (adapter != tbl_adapter) { build_table2<NR>(L, tbl, s_code, aoff, n, r0, lane, s_match, s_mis, b); tbl_adapter = adapter; }

        uint32_t ex[2] = {0u, 0u}, ey[2] = {0u, 0u};
#pragma unroll 1
        for (int sub = 0; sub < 2; ++sub) {
            const HalfTask t = tk[sub];
            const int len = t.len;
            const int mw = __reduce_max_sync(0xffffffffu, len);
            if (mw == 0) {                                     // nobody has a task in this half
                if (t.out_idx >= 0) { if (TRACE) write_failed(L.out, t.out_idx); else L.scan_out[t.out_idx] = (int)0x80000000; }
                continue;
            }
            const bool fin = len > 0 && t.col0 + len == t.m_true;
            int lo1 = t.own0 + 1, hi1 = fin ? len - 1 : len;   // candidate columns of the last adapter row
            if (t.forced >= 0) { if (t.forced <= hi1) { lo1 = max(lo1, t.forced); hi1 = t.forced; } else hi1 = -1; }
            const bool act = t.out_idx >= 0 && len > 0;
            const int ulo = __reduce_max_sync(0xffffffffu, act ? lo1 : 0);
            const int ufin = __reduce_min_sync(0xffffffffu, fin ? len : 0x7fffffff);
            const int uhi = min(__reduce_min_sync(0xffffffffu, act ? hi1 : 0x7fffffff), ufin - 2);
            const bool z = len > 0 && t.col0 == 0 && t.own0 == 0 && (t.forced < 0 || t.forced == 0);
            uint32_t bestS = pk(0, z ? OFF : 0);               // the last adapter row lives in the high half
            int ord = 0, finS = 0, finI = 0, finE = 0;
            const uint8_t *const seq = L.arena + t.off;
            const int lim = max(len - 1, 0);

            uint32_t Mo[NR], Hl[NR];
#pragma unroll
            for (int r = 0; r < NR; ++r) { Mo[r] = pk(B0 + b * (r + 1), B0 + b * (NR + r + 1)); Hl[r] = 0; }
            uint32_t rowT = pk2(B0);
            uint32_t bM1 = pk2(B0 + b * NR), bM2 = bM1, bV1 = 0;     // row NR at columns t-1, t-2 (high halves count)
            const uint32_t candK = pk2(OFF - b * R2 + b);     // candidate = M(n, t-1) + candK - rowT(t) = score + OFF
            uint32_t *const ckp = TRACE ? L.scratch + (uint64_t)(unit - L.unit0) * 32u * ckw + (uint64_t)sub * 16u * ckw + (uint64_t)lane * 8u : nullptr;
            uint32_t ckoff = 0;
            int ckc = 0;
            Raw4 nx = load4(seq, 0, lim);
            uint32_t prev = 0;

            auto column = [&](const int j, const uint32_t c_lo, const uint32_t c_hi, const bool general) {
                const uint32_t *tp = tbl + pair_index(c_lo, c_hi);
                const uint32_t rowTp = rowT;
                rowT += Bp;
                const uint32_t vb = __viaddmax_s16x2(bV1, GEB, bM1);       // V(NR + 1, t - 1) from the row above
                uint32_t Ml, Vl;
                col_score2<NR>(Mo, Hl, tp, lo_hi(rowTp, bM2), lo_hi(rowT, vb), GEB, Ml, Vl);
                bM2 = bM1; bM1 = swap16(Ml); bV1 = swap16(Vl);
                uint32_t cand = (Ml + candK - rowT) & 0xFFFF0000u;
                if (general) { if (!(j - 1 >= lo1 && j - 1 <= hi1)) cand = 0u; }
                bool keep_hi, keep_lo;
                bestS = vmax_keep(bestS, cand, keep_hi, keep_lo);
                if (!keep_hi) ord = j - 1;
                if (general) {
                    if (j == 1) {                                // the high half has not started yet: back to its column 0
#pragma unroll
                        for (int r = 0; r < NR; ++r) { Mo[r] = (Mo[r] & 0xFFFFu) | ((uint32_t)(B0 + b * (NR + r + 1)) << 16); Hl[r] &= 0xFFFFu; }
                    }
                    if (fin && j == len) {                       // final column, rows of the first stripe
#pragma unroll
                        for (int r = 0; r < NR; ++r) { s_fl[(2 * r) * PCB_BLOCK] = Mo[r]; s_fl[(2 * r + 1) * PCB_BLOCK] = Hl[r]; }
                    }
                    if (fin && j == len + 1) {                   // ... and of the second, one iteration later
#pragma unroll
                        for (int r = 0; r < NR; ++r) {
                            s_fl[(2 * r) * PCB_BLOCK] = lo_hi(s_fl[(2 * r) * PCB_BLOCK], Mo[r]);
                            s_fl[(2 * r + 1) * PCB_BLOCK] = lo_hi(s_fl[(2 * r + 1) * PCB_BLOCK], Hl[r]);
                        }
                        // cells (i, len), i = 1..n, in order; V rebuilt as the forward pass computed it (as in k_fast)
                        int Tn = lo16(rowT) - b + b * r0, Vp = 0, Mp = Tn, best = 0, bi = 0, be = 0;     // rowT: bias of row 0 at column len + 1
                        for (int i = 1; i <= n; ++i) {
                            const int rr = r0 + i - 1, h = rr >= NR ? 1 : 0, r = rr - NR * h;
                            const int M = half16(s_fl[(2 * r) * PCB_BLOCK], h), H = half16(s_fl[(2 * r + 1) * PCB_BLOCK], h);
                            const int V = max(Vp + geb, Mp);
                            Tn += b;
                            const int c = M + OFF - Tn;
                            if ((t.forced < 0 || t.forced == len + i) && c > best) { best = c; bi = i; be = ((V == M) ? 2 : 0) | ((H == M) ? 1 : 0); }
                            Vp = V; Mp = M;
                        }
                        finS = best; finI = bi; finE = be;
                    }
                }
            };

            for (int j0 = 1; j0 <= mw + 1; j0 += 4) {
                const uint32_t cur = join4(nx);
                nx = load4(seq, j0 + 3, lim);
                uint32_t c_lo = cur, c_hi = __funnelshift_l(prev, cur, 8);      // the high half is one column behind
                prev = cur;
                if (TRACE && ckc == 0 && j0 > 1) {               // checkpoint k = state after iteration j0 - 1 = k * C
                    uint32_t *const cq = ckp + ckoff;
                    PCF_CHECK(L, (uint64_t)ckoff + (G8 + 1) * 256u <= 16u * ckw);
                    const uint32_t T1 = GEB + 0x00010001u;
#pragma unroll
                    for (int g = 0; g < G8; ++g) {
                        uint32_t v[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) { const int r = 8 * g + e; v[e] = r < NR ? Mo[r < NR ? r : 0] + __vminu2(Mo[r < NR ? r : 0] - Hl[r < NR ? r : 0], T1) * 8192u : 0u; }
                        st256(cq + g * 256, v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7]);
                    }
                    st256(cq + G8 * 256, bM1, bM2, bV1, 0u, 0u, 0u, 0u, 0u);
                    ckoff += (G8 + 1) * 256;
                }
                if (TRACE) { ckc += 4; if (ckc == C) ckc = 0; }
                if (!TRACE && ((j0 - 1) & (REBASE - 1)) == 0 && j0 > 1 && !(fin && j0 - 1 == len)) {   // keep the growing bias inside 15 bits (not between the two halves of a final-column dump)
                    const uint32_t d = pk2(b * REBASE);
#pragma unroll
                    for (int r = 0; r < NR; ++r) { Mo[r] -= d; Hl[r] -= d; }
                    rowT -= d; bM1 -= d; bM2 -= d; bV1 -= d;
                }
                if (j0 - 1 >= ulo && j0 + 2 <= uhi && j0 > 1) {
#pragma unroll UNR
                    for (int k = 0; k < 4; ++k) { column(j0 + k, c_lo & 0xFFu, c_hi & 0xFFu, false); c_lo >>= 8; c_hi >>= 8; }
                } else {
#pragma unroll 1
                    for (int k = 0; k < 4; ++k) { column(j0 + k, c_lo & 0xFFu, c_hi & 0xFFu, true); c_lo >>= 8; c_hi >>= 8; }
                }
            }

            int sc = (int)(bestS >> 16);
            bool fc = false;
            if (finS > sc) { sc = finS; ord = len + finI; fc = true; }
            if (!TRACE) {
                if (t.out_idx >= 0) L.scan_out[t.out_idx] = (len > 0 && sc > 0) ? (int)(((uint32_t)(sc - OFF) << 16) | ((uint32_t)(0x3FFF - ord) << 2)) : (int)0x80000000;
                continue;
            }
            if (t.out_idx < 0) continue;
            if (len <= 0) { write_failed(L.out, t.out_idx); continue; }
            if (sc <= 0) { atomicOr(L.err_flag, 2); write_failed(L.out, t.out_idx); continue; }
            int pci, pcj, pst;
            if (fc) { pci = ord - len; pcj = len; pst = (finE & 2) ? 1 : ((finE & 1) ? 2 : 0); }
            else    { pci = n; pcj = ord; pst = 4; }
            ex[sub] = (uint32_t)pci | ((uint32_t)pcj << 7) | ((uint32_t)pst << 17) | 0x80000000u;
            ey[sub] = (uint32_t)sc;
        }
        if (TRACE) *entp = make_uint4(ex[0], ey[0], ex[1], ey[1]);
    }
}

// One op of the reverse walk of the two-stripe kernels: arg = register row 0 .. 2*NR-1 of the walk's adapter row, jcol = its
// column - (first column of the block).  The flags of stripe h, row arg - NR*h, column jcol are in iteration slot jcol + h.
template <int NR>
__device__ __forceinline__ void walk_step2(const uint32_t *s_fl, int &arg, int &jcol, uint32_t &pos2, Walk &w)
{
    constexpr int FW = NR / 4;
    const int h = arg >= NR ? 1 : 0, ar = arg - NR * h;
#ifdef PCF_CHECKED
    if (ar < 0 || (ar >> 2) >= FW || jcol + h < 0 || jcol + h >= 16) asm volatile("trap;");
#endif
    const uint32_t f = walk_flags<FW>(s_fl, h, ar, jcol + h);
    const uint32_t fh = f & 1u, fg = (f >> 1) & 1u, fv = (f >> 16) & 1u, fd = (f >> 17) & 1u;
    const uint32_t sM = w.s == 0 ? 1u : 0u, sV = w.s == 1 ? 1u : 0u;
    const uint32_t isD = sM & fd, isV = sV | (sM & (fd ^ 1u) & fg), isH = (isD | isV) ^ 1u;
    const uint32_t extV = isV & fv, extH = isH & fh;
    walk_acc_step(w.acc, pos2, isD, isV, isH, extV | extH);
    w.s = (int)(extV | (extH << 1));
    arg -= (int)(isH ^ 1u); jcol -= (int)(isV ^ 1u);
}

// k_rounds2<NR, C>: pass 2 of the two-stripe trace.  Work items, sorting, generations and results as k_rounds; a lane takes
// ONE item at a time (its two stripes fill the packed halves), so a chunk of 64 items is two sub-rounds of 32.
// A block = iterations k*C+1 .. k*C+C: columns k*C+1.. of the first stripe, k*C.. of the second.
// dynamic smem: [W][NR*25] score tables, [W][2*NR] adapter codes, then per thread C*NR/4 flag words.
template <int NR, int C>
__global__ void __launch_bounds__(PCB_BLOCK, 1) k_rounds2(const PcbLaunch L)
{
    extern __shared__ uint32_t smem[];
    constexpr int FW = NR / 4, W = PCB_BLOCK / 32, R2 = 2 * NR, G8 = (NR + 7) / 8;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *const tbl = smem + warp * (NR * TP);
    uint32_t *const s_code = smem + W * (NR * TP) + warp * R2;
    uint32_t *const s_fl = smem + W * (NR * TP + R2) + threadIdx.x;

    const Meta M = meta_of(L);
    const int A = L.n_major;
    const bool finish = L.finish != 0;
    const int b = -L.sc.gap_open;
    const int s_match = L.sc.match + 2 * b, s_mis = L.sc.mismatch + 2 * b;
    const uint32_t GEB = pk2(L.sc.gap_extend + b), Bp = pk2(b);
    const bool cross = L.src.units == nullptr;
    const int64_t groups = cross ? (L.src.n_windows + 63) / 64 : 0;
    const uint64_t ckw = ckw2_of(L.mmax, C, NR);

    // Work is handed out 32 items (half a chunk) at a time from a queue head in meta, the LAST chunks first: inside a major
    // the sort puts the longest paths last, and a round costs its longest path, so the expensive rounds start early and the
    // cheap ones fill the tail.  (A static split of the chunk list left SMs idle for half of the kernel: some warps drew
    // nothing but 13-block rounds.)
    const int64_t total = M.chunk_base[A];
    int32_t *const qhead = M.counters + (L.gen < 3 ? 1 : 2) + L.gen;
    int a = -1, n = 0, r0 = 0;
    int32_t tbl_adapter = -1, out_col = 0;
    int64_t a_item0 = 0, a_item1 = 0, a_chunk0 = 0;

    for (;;) {
        int64_t q = 0;
        if (lane == 0) q = atomicAdd(qhead, 1);
        q = __shfl_sync(0xffffffffu, q, 0);
        if (q >= 2 * total) break;
        const int64_t qq = 2 * total - 1 - q, c = qq >> 1;
        const int sub = (int)(qq & 1);
        if (a < 0 || c < a_chunk0 || c >= M.chunk_base[a + 1]) {     // another major: the one whose chunk range holds c
            int lo = 0, hi = A - 1;
            while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (M.chunk_base[mid] <= c) lo = mid; else hi = mid - 1; }
            a = lo;
            a_item0 = M.base[(int64_t)a * NB]; a_item1 = M.base[(int64_t)(a + 1) * NB]; a_chunk0 = M.chunk_base[a];
            const int32_t adapter = cross ? L.src.ids[a] : a;
            if (cross) out_col = L.src.id_pos ? L.src.id_pos[a] : a;
            const int64_t aoff = L.ad.off[adapter];
            n = (int)(L.ad.off[adapter + 1] - aoff);
            r0 = R2 - n;
            if (adapter != tbl_adapter) { build_table2<NR>(L, tbl, s_code, aoff, n, r0, lane, s_match, s_mis, b); tbl_adapter = adapter; }
        }
        {
            const int64_t i0 = a_item0 + (c - a_chunk0) * 64 + 32 * sub + lane;
            bool valid = i0 < a_item1;
            uint32_t ref, endc;
            Walk wk;
            uint64_t off;
            {
                uint32_t v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                if (valid) ld256(reinterpret_cast<const uint32_t *>(L.sorted + i0), v);
                ref = v[0]; endc = v[2];
                walk_unpack(wk, v[1], v[3], v[4], v[5]);
                off = (uint64_t)v[6] | ((uint64_t)v[7] << 32);
            }
            if (!__any_sync(0xffffffffu, valid)) continue;
            const uint8_t *const seq = L.arena + off;
            constexpr int lim = 0x7fffffff;
            // this task's checkpoints: written by lane (slot >> 1) as its task (slot & 1)
            const uint32_t *const ckp = L.scratch + ((uint64_t)(ref >> 6) * 32u * ckw + (uint64_t)(ref & 1u) * 16u * ckw + (uint64_t)((ref & 63u) >> 1) * 8u);

            for (;;) {
                const bool go = valid && wk.i > 0 && wk.j > 0;
                const int arg0 = r0 + wk.i - 1, h0 = arg0 >= NR ? 1 : 0;
                const int k = go ? (wk.j + h0 - 1) / C : 0, jb = k * C;
                const int tmax = __reduce_max_sync(0xffffffffu, go ? wk.j + h0 - jb : 0);
                const int rlim = __reduce_max_sync(0xffffffffu, go ? (h0 ? NR : ((arg0 + 4) & ~3)) : 0);
                if (tmax > 0) {
                    PCF_CHECK(L, tmax <= C && rlim <= NR);
                    // the block's read bases: low stripe = columns jb+1.., high stripe one column behind
                    uint32_t ql[C / 4], qh[C / 4];
                    {
                        const uint32_t before = k > 0 ? join4(load4(seq, jb - 4, lim)) : 0u;
#pragma unroll
                        for (int g = 0; g < C / 4; ++g) ql[g] = join4(load4(seq, jb + 4 * g, lim));
#pragma unroll
                        for (int g = 0; g < C / 4; ++g) qh[g] = __funnelshift_l(g ? ql[g > 0 ? g - 1 : 0] : before, ql[g], 8);
                    }
                    uint32_t Mo[NR], Hl[NR];
                    uint32_t bM1 = pk2(B0 + b * NR), bM2 = bM1, bV1 = 0u;
                    {
                        const bool ld = go && k > 0;
                        const uint32_t *const p = ckp + (uint64_t)max(k - 1, 0) * (G8 + 1) * 256u;
                        const uint32_t T1s = (GEB + 0x00010001u) * 8192u;
                        PCF_CHECK(L, !ld || ((int64_t)(ref >> 6) < L.unit1 - L.unit0 && (uint64_t)k * (G8 + 1) * 16u <= ckw));
#pragma unroll
                        for (int g = 0; g < G8; ++g) {
                            if (8 * g < rlim) {
                                uint32_t x[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e) x[e] = pk(B0 + b * (8 * g + e + 1), B0 + b * (NR + 8 * g + e + 1)) + T1s;
                                if (ld) ld256(p + g * 256, x);
#pragma unroll
                                for (int e = 0; e < 8; ++e) {
                                    if (8 * g + e < NR) {
                                        const uint32_t m = x[e] & 0x1FFF1FFFu;
                                        Mo[8 * g + e] = m; Hl[8 * g + e] = m - ((x[e] >> 13) & 0x00070007u);
                                    }
                                }
                            } else {
#pragma unroll
                                for (int e = 0; e < 8; ++e) if (8 * g + e < NR) { Mo[8 * g + e] = 0u; Hl[8 * g + e] = 0u; }
                            }
                        }
                        if (ld) { uint32_t x[8]; ld256(p + G8 * 256, x); bM1 = x[0]; bM2 = x[1]; bV1 = x[2]; }
                    }
                    uint32_t rT = pk2(B0 + b * jb);
                    const bool pend = go && (wk.s & 4);
                    const bool anypend = __any_sync(0xffffffffu, pend);
                    int s = wk.s & 3;
#pragma unroll 1
                    for (int t = 1; t <= tmax; ++t) {
                        const uint32_t *tp = tbl + pair_index(ql[0] & 0xFFu, qh[0] & 0xFFu);
#pragma unroll
                        for (int g = 0; g < C / 4; ++g) {                  // shift the base queues by one byte
                            ql[g] = __funnelshift_r(ql[g], g + 1 < C / 4 ? ql[(g + 1) % (C / 4)] : 0u, 8);
                            qh[g] = __funnelshift_r(qh[g], g + 1 < C / 4 ? qh[(g + 1) % (C / 4)] : 0u, 8);
                        }
                        const uint32_t rowTp = rT;
                        rT += Bp;
                        uint32_t Ml, Vl, Hlast;
                        col_flags2<NR>(Mo, Hl, tp, lo_hi(rowTp, bM2), lo_hi(rT, bM1), bV1 & 0xFFFF0000u, GEB,
                                       s_fl + (t - 1) * FW * PCB_BLOCK, rlim, Ml, Vl, Hlast);
                        bM2 = bM1; bM1 = swap16(Ml); bV1 = swap16(Vl);
                        if (k == 0 && t == 1) {                            // the high stripe starts from its column 0 next iteration
#pragma unroll
                            for (int r = 0; r < NR; ++r) { Mo[r] = (Mo[r] & 0xFFFFu) | ((uint32_t)(B0 + b * (NR + r + 1)) << 16); Hl[r] &= 0xFFFFu; }
                        }
                        if (anypend) {
                            // end cell in the last adapter row (second stripe, last register row): V == M / H == M decide how
                            // the trace starts (_correctTraceValue, dp_algorithm_impl.h:1169-1185)
                            if (pend && jb + t == wk.j + 1) s = ((Ml ^ Vl) >> 16) == 0 ? 1 : ((((Ml ^ Hlast) >> 16) == 0) ? 2 : 0);
                        }
                    }
                    __syncwarp();
                    wk.s = s;                                              // the pending bit is spent
                    {
                        int arg = arg0, jcol = wk.j - jb - 1;
                        uint32_t pz = (uint32_t)(2 * wk.j + 1) | ((uint32_t)(2 * wk.i + 1) << 16);
                        bool act = go;
                        while (act) {
                            walk_step2<NR>(s_fl, arg, jcol, pz, wk);
                            act = arg >= r0 && jcol + jb >= 0 && jcol + (arg >= NR ? 1 : 0) >= 0;
                        }
                        if (go) { wk.i = arg - r0 + 1; wk.j = jcol + jb + 1; }
                    }
                    __syncwarp();                                          // the flag words are rewritten by the next round
                }

                // paths that reached row 0 / column 0 are complete (this also covers the empty path)
                const bool keep = valid && wk.i > 0 && wk.j > 0;
                if (valid && !keep) {
                    valid = false;
                    int col0 = 0, mtrue;
                    int64_t oidx;
                    if (cross) {
                        const int64_t wi = (L.unit0 + (int64_t)(ref >> 6) - (int64_t)a * groups) * 64 + (int64_t)(ref & 63u);
                        mtrue = L.src.win_len[wi]; oidx = wi * L.src.n_ids_total + out_col;
                    } else {
                        const PcbTask &tk = L.src.tasks[L.src.units[L.unit0 + (int64_t)(ref >> 6)].task0 + (int64_t)(ref & 63u)];
                        col0 = tk.col0; mtrue = tk.m_true; oidx = tk.out_idx;
                    }
                    const int endi = (int)(endc & 127u), endj = (int)((endc >> 7) & 0x3FFu), score = (int)(endc >> 17) - OFF;
                    if (wk.j == 0 && wk.i > 0 && col0 > 0) atomicOr(L.err_flag, 1);      // window too narrow for this path
                    int ok;
                    const int matches = walk_acc_matches(wk.acc, endi, endj, wk.i, wk.j, score, L.sc.match, L.sc.mismatch, L.sc.gap_open, L.sc.gap_extend, &ok);
                    if (!ok) atomicOr(L.err_flag, 16);
                    PathAcc pa;
                    walk_acc_finish(wk.acc, endi, endj, wk.i, wk.j, matches, pa);
                    PathStats ps;
                    path_finalize(pa, (int64_t)col0 + wk.j, wk.i, (int64_t)mtrue - (col0 + endj), n - endi, mtrue, n, ps);
                    pcb_result res;
                    res.read_start = ps.read_start; res.read_end = ps.read_end;
                    res.adapter_start = ps.adapter_start; res.adapter_end = ps.adapter_end;
                    res.raw_score = score; res.matches = ps.matches;
                    res.aligned_len = ps.aligned_len; res.full_len = ps.full_len;
                    PCF_CHECK(L, oidx >= 0 && (L.n_out <= 0 || oidx < L.n_out));
                    L.out[oidx] = res;
                }
                if (finish) {
                    if (!__any_sync(0xffffffffu, keep)) break;
                    continue;                                              // these 32 stay together until the last is done
                }
                // the others go to the next generation, counted per (major, what is left of the path); counters as in k_rounds
                int32_t *const n_surv = M.counters + ((L.gen & 1) ? 4 : 0), *const hist_next = (L.gen & 1) ? M.hist0 : M.hist1;
                const uint32_t m0 = __ballot_sync(0xffffffffu, keep);
                const int nk = __popc(m0);
                if (nk) {
                    int first = 0;
                    if (lane == 0) first = atomicAdd(n_surv, nk);
                    first = __shfl_sync(0xffffffffu, first, 0);
                    PCF_CHECK(L, first + nk <= 64 * (L.unit1 - L.unit0));
                    const uint32_t pos = walk_pos(wk);
                    const int key = keep ? a * NB + bucket1_two(pos) : -1;
                    const uint32_t mk = __match_any_sync(0xffffffffu, key);
                    if (keep) {
                        if (lane == __ffs(mk) - 1) atomicAdd(hist_next + key, __popc(mk));
                        put_item(L.surv + first + __popc(m0 & ((1u << lane) - 1u)), ref, pos, endc, wk.acc.kAR, wk.acc.tHV, wk.acc.nHV, off);
                    }
                }
                break;
            }
        }
    }
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host-side launchers, by kernel class
// ------------------------------------------------------------------------------------------------
static size_t fast_smem(int kind, int nr, int c, bool two = false)
{
    const size_t W = PCB_BLOCK / 32;
    if (two)             // two-stripe kernels: codes of 2 * nr rows; flag words or column dump as below
        return (W * (size_t)(nr * TP + 2 * nr) + (size_t)PCB_BLOCK * (kind == 2 ? (size_t)c * (nr / 4) : (size_t)2 * nr)) * sizeof(uint32_t);
    if (kind == 2)       // rounds: tables, codes, then the flag words per thread
        return (W * (size_t)(nr * TP + nr) + (size_t)PCB_BLOCK * ((size_t)c * (nr / 4))) * sizeof(uint32_t);
    return (W * (size_t)(nr * TP + nr) + (size_t)PCB_BLOCK * 2 * nr) * sizeof(uint32_t);      // column dump
}

template <int MODE, int NR, bool EXACT, int C>
static cudaError_t launch_fast(const PcbLaunch &L, int grid, cudaStream_t st)
{
    static bool attr = false;
    const size_t sm = fast_smem(MODE, NR, C);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(k_fast<MODE, NR, EXACT, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    k_fast<MODE, NR, EXACT, C><<<grid, PCB_BLOCK, sm, st>>>(L);
    return cudaGetLastError();
}

template <int NR, bool EXACT, int C>
static cudaError_t launch_rounds(const PcbLaunch &L, int grid, cudaStream_t st)
{
    static bool attr = false;
    const size_t sm = fast_smem(2, NR, C);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(k_rounds<NR, EXACT, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    k_rounds<NR, EXACT, C><<<grid, PCB_BLOCK, sm, st>>>(L);
    return cudaGetLastError();
}

// the two-stripe classes: 56 or 72 register rows x 2 stripes
constexpr int C2 = PCF_C_LARGE;
template <int MODE, int NR>
static cudaError_t launch_fast2(const PcbLaunch &L, int grid, cudaStream_t st)
{
    static bool attr = false;
    const size_t sm = fast_smem(MODE, NR, C2, true);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(k_fast2<MODE, NR, C2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    k_fast2<MODE, NR, C2><<<grid, PCB_BLOCK, sm, st>>>(L);
    return cudaGetLastError();
}
template <int NR>
static cudaError_t launch_rounds2(const PcbLaunch &L, int grid, cudaStream_t st)
{
    static bool attr = false;
    const size_t sm = fast_smem(2, NR, C2, true);
    if (!attr) {
        cudaError_t e = cudaFuncSetAttribute(k_rounds2<NR, C2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        if (e != cudaSuccess) return e;
        attr = true;
    }
    k_rounds2<NR, C2><<<grid, PCB_BLOCK, sm, st>>>(L);
    return cudaGetLastError();
}
template <int KIND, int NR>
static int occ_two()
{
    int per = 0;
    const size_t sm = fast_smem(KIND, NR, C2, true);
    if (KIND == 2) {
        cudaFuncSetAttribute(k_rounds2<NR, C2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_rounds2<NR, C2>, PCB_BLOCK, sm);
    } else {
        cudaFuncSetAttribute(k_fast2<(KIND == 1), NR, C2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_fast2<(KIND == 1), NR, C2>, PCB_BLOCK, sm);
    }
    return per < 1 ? 1 : per;
}
#define PCF_BY_TWO(...) do { if (cls == 6) { constexpr int NR = 56; __VA_ARGS__; } else { constexpr int NR = 72; __VA_ARGS__; } } while (0)

template <int KIND, int NR, bool EXACT, int C>
static int occ_one()
{
    int per = 0;
    const size_t sm = fast_smem(KIND, NR, C);
    if (KIND == 2) {
        cudaFuncSetAttribute(k_rounds<NR, EXACT, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_rounds<NR, EXACT, C>, PCB_BLOCK, sm);
    } else {
        cudaFuncSetAttribute(k_fast<(KIND == 1), NR, EXACT, C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per, k_fast<(KIND == 1), NR, EXACT, C>, PCB_BLOCK, sm);
    }
    return per < 1 ? 1 : per;
}

// kernel classes: see pcb_kernels.h
#define PCF_BY_CLASS(...)                                                                 \
    switch (cls) {                                                                          \
    case 0:  { constexpr int NR = 24; constexpr bool EX = true;  constexpr int C = PCF_C_SMALL; __VA_ARGS__; } break; \
    case 1:  { constexpr int NR = 24; constexpr bool EX = false; constexpr int C = PCF_C_SMALL; __VA_ARGS__; } break; \
    case 2:  { constexpr int NR = 32; constexpr bool EX = false; constexpr int C = PCF_C_SMALL; __VA_ARGS__; } break; \
    case 3:  { constexpr int NR = 40; constexpr bool EX = false; constexpr int C = PCF_C_LARGE; __VA_ARGS__; } break; \
    case 4:  { constexpr int NR = 52; constexpr bool EX = false; constexpr int C = PCF_C_LARGE; __VA_ARGS__; } break; \
    default: { constexpr int NR = 72; constexpr bool EX = false; constexpr int C = PCF_C_LARGE; __VA_ARGS__; } break; \
    }

cudaError_t pcf_launch_scan(const PcbLaunch &L, int cls, cudaStream_t st)
{
    if (L.unit1 <= L.unit0) return cudaSuccess;
    if (cls < 0 || cls >= PCF_NCLS) return cudaErrorInvalidValue;
    if (pcf_class_two(cls)) { PCF_BY_TWO(return (launch_fast2<0, NR>(L, L.grid, st))); }
    PCF_BY_CLASS(return (launch_fast<0, NR, EX, C>(L, L.grid, st)))
    return cudaSuccess;
}

// Trace of the units [L.unit0, L.unit1), in two parts: pass 1 ...
cudaError_t pcf_launch_pass1(const PcbLaunch &L, int cls, cudaStream_t st)
{
    if (L.unit1 <= L.unit0) return cudaSuccess;
    if (cls < 0 || cls >= PCF_NCLS) return cudaErrorInvalidValue;
    cudaError_t e = cudaSuccess;
    if (pcf_class_two(cls)) { PCF_BY_TWO(return (launch_fast2<1, NR>(L, L.grid, st))); }
    PCF_BY_CLASS(e = launch_fast<1, NR, EX, C>(L, L.grid, st))
    return e;
}

// pass 2: per generation count / scan / scatter / rounds.  Nothing comes back to the host.
cudaError_t pcf_launch_pass2(const PcbLaunch &L0, int cls, int grid_rounds, cudaStream_t st)
{
    if (L0.unit1 <= L0.unit0) return cudaSuccess;
    if (cls < 0 || cls >= PCF_NCLS) return cudaErrorInvalidValue;
    PcbLaunch L = L0;
    const int C = pcf_class_c(cls), two = pcf_class_two(cls);
    const int64_t nu = L.unit1 - L.unit0;
    cudaError_t e = cudaMemsetAsync(L.meta, 0, pcf_meta_words(L.n_major) * sizeof(int32_t), st);
    if (e != cudaSuccess) return e;
    const unsigned g0 = (unsigned)((nu * 32 + 255) / 256);
    k_count0<<<g0, 256, 0, st>>>(L, C, two);
    k_bucket_scan<<<1, 1024, 0, st>>>(L, 0);
    k_scatter0<<<g0, 256, 0, st>>>(L, C, two);
    const int gens = pcf_generations(cls);
    auto rounds = [&](int gen) -> cudaError_t {
        L.gen = gen; L.finish = gen == gens - 1;
        cudaError_t r = cudaSuccess;
        if (two) PCF_BY_TWO(r = launch_rounds2<NR>(L, grid_rounds, st));
        else { PCF_BY_CLASS(r = launch_rounds<NR, EX, C>(L, grid_rounds, st)) }
        return r;
    };
    if ((e = rounds(0)) != cudaSuccess) return e;
    // the survivors are counted on the device: the scatters are sized for the worst case and leave early
    const unsigned g1 = (unsigned)((nu * 64 + 255) / 256);
    k_bucket_scan<<<1, 1024, 0, st>>>(L, 1);
    k_scatter1<<<g1, 256, 0, st>>>(L, two, 0);
    if ((e = rounds(1)) != cudaSuccess) return e;
    for (int g = 2; g < gens; ++g) {        // later blocks one by one too: what is left after each is sorted again
        k_bucket_scan<<<1, 1024, 0, st>>>(L, (g - 1) & 1 ? 0 : 1);
        k_scatter1<<<g1, 256, 0, st>>>(L, two, (g - 1) & 1 ? 4 : 0);
        if ((e = rounds(g)) != cudaSuccess) return e;
    }
    return cudaGetLastError();
}

// generations of pass 2: the two-stripe classes take two (first block, then the rest); the others PCB_GENS (default below)
int pcf_generations(int cls)
{
    static const int env = [] { const char *g = getenv("PCB_GENS"); return g ? atoi(g) : 0; }();
    static const int env_small = [] { const char *g = getenv("PCB_GENS_SMALL"); return g ? atoi(g) : 0; }();
    static const int env_two = [] { const char *g = getenv("PCB_GENS_TWO"); return g ? atoi(g) : 0; }();
    if (pcf_class_two(cls)) return env_two >= 2 && env_two <= 5 ? env_two : 2;
    // measured (131,072 windows x 108 24-base sequences): 2 / 3 / 4 / 5 generations = 13.08 / 12.58 / 12.44 / 12.43 ms; 16,384 windows
    // x 12 68-base sets: 0.850 / 0.817 / 0.844 / 0.865 ms; the small classes lose with a third (more launches than items)
    if (cls == 0 || cls == 5) return env >= 2 && env <= 5 ? env : (cls == 0 ? 4 : 3);
    return env_small >= 2 && env_small <= 5 ? env_small : 2;
}
int pcf_trace_launches(int cls) { return 5 + 3 * (pcf_generations(cls) - 1); }

int pcf_occupancy_grid(int kind, int cls)
{
    int dev = 0, sms = 0, per = 1;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (pcf_class_two(cls)) PCF_BY_TWO(per = kind == 0 ? occ_two<0, NR>() : (kind == 1 ? occ_two<1, NR>() : occ_two<2, NR>()));
    else if (kind == 0) { PCF_BY_CLASS(per = (occ_one<0, NR, EX, C>())) }
    else if (kind == 1) { PCF_BY_CLASS(per = (occ_one<1, NR, EX, C>())) }
    else                { PCF_BY_CLASS(per = (occ_one<2, NR, EX, C>())) }
    return sms * per;
}
