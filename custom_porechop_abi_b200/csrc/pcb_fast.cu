/*
 * pcb_fast.cu -- the fast sm_100a kernels of the adapter-alignment hot path (affine schemes, ACGT adapters of <= 52 bases).
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
 *      actually walks through (1.8 of them on average), keeps the flags in shared memory and walks them.  Values are
 *      15-bit non-negative, so "a >= b" is bit 15 of (a + 0x8000 - b): one IADD3 per flag, extracted four at a time with
 *      PRMT in sign-replicate mode.  Two kernels because the passes want different things: pass 1 is a register-bound
 *      integer stream with no memory stalls, the rounds are short latency-bound phases (checkpoint fetch, flag walk) that
 *      need flag memory and resident warps; fused into one kernel (round 1) each held the other back.
 *   3. the scan kernel for long reads (middle-adapter search) is pass 1 without checkpoints (k_fast<0>), with the bias
 *      re-based every 1024 columns.
 *   4. no copies, no waits.  The column body is straight-line code over NR rows (template parameter: 24 = the length of
 *      almost every Porechop adapter and barcode, exact; 24 / 32 / 40 / 52 = any shorter-or-equal adapter, bottom-aligned,
 *      the rows above the adapter being exact copies of DP row 0), so ptxas updates the DP registers in place; read bases
 *      are fetched four columns at a time, one group ahead of their use, so that no column waits on a global load.
 *
 * One thread = two tasks (the halves of each 32-bit register), one warp = 64 tasks sharing an adapter, as in pcb_kernels.cu.
 * Sequence codes in the arenas are Dna5 code * 8 (0, 8, 16, 24, 32 = N), so that a code is directly a byte shift.
 */
#include "pcb_kernels.h"
#include "align_stats.h"

#include <cuda_runtime.h>
#include <stdint.h>

namespace {

// CTAs per SM the register allocation is held to.  Measured on B200: the score-only kernels issue integer instructions at the
// pipes' limit with 2-3 CTAs per SM already; a larger register budget (no spills, fewer register moves) beats more warps.
#ifndef PCF_MINB_SCAN
#define PCF_MINB_SCAN 2
#endif
#ifndef PCF_MINB_P1
#define PCF_MINB_P1(NR) ((NR) <= 32 ? 3 : 2)
#endif
#ifndef PCF_MINB_P2
#define PCF_MINB_P2(NR) ((NR) <= 24 ? 3 : 2)      /* shared memory (flags + queue) allows no more */
#endif
#define PCF_MINB(MODE, NR) ((MODE) ? PCF_MINB_P1(NR) : PCF_MINB_SCAN)
#ifndef PCF_UNROLL
#define PCF_UNROLL 2
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
// column, stride PCB_BLOCK.
template <int NR, bool EXACT>
__device__ __forceinline__ void col_flags(uint32_t (&Mo)[NR], uint32_t (&Hl)[NR], const uint32_t *tp,
                                          uint32_t rowTp, uint32_t rowT, uint32_t GEB, uint32_t *fp,
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

__device__ __forceinline__ uint64_t ckw_of(int mmax, int C, int NR) { return (uint64_t)((mmax + C - 1) / C + 1) * 2ull * NR; }   // = pcf_ck_words (pcb_kernels.h)

// One op of the reverse walk for half h of this thread (dp_traceback_impl.h:379-451).  Flag word layout as col_flags writes
// it; the block's read bases sit behind the C*FW flag words.  State 0 = M, 1 = V, 2 = H.  In state M a cell whose diagonal
// flag is off hands over to V or H and that state's step is taken at once, from the same cell's flags, as SeqAn does.
template <int FW, int C>
__device__ __forceinline__ void walk_step(const uint32_t *s_fl, const uint32_t *s_code, const int h, const int base, const int r0,
                                          int &i, int &j, int &s, WalkAcc &a)
{
    const int ar = r0 + i - 1, jc = j - base - 1;         // register row of adapter row i
    const uint32_t f = s_fl[(jc * FW + (ar >> 2)) * PCB_BLOCK] >> (8 * h + 2 * (ar & 3));
    const bool fh = f & 1u, fg = f & 2u, fv = f & 0x10000u, fd = f & 0x20000u;
    if (s == 0 && fd) {
        const uint32_t code8 = (s_fl[(C * FW + h * (C / 4) + (jc >> 2)) * PCB_BLOCK] >> (8 * (jc & 3))) & 0xFFu;
        walk_acc_D(a, i, j, code8 == s_code[ar] * PCB_CODE_SCALE);
        --i; --j;
    } else if (s == 1 || (s == 0 && fg)) {
        walk_acc_V(a, i, j);
        --i; s = fv ? 1 : 0;
    } else {
        walk_acc_H(a, i, j);
        --j; s = fh ? 2 : 0;
    }
}

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
// Entry of a task (two words): x = end row | end column << 6 | start state << 18 | 1 << 31 (0: no entry, the failed record
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
        // this thread's checkpoint records of the unit: [unit - unit0][lane][k][2 * NR]
        uint32_t *const ckp = TRACE ? L.scratch + ((uint64_t)(unit - L.unit0) * 32u + (uint64_t)lane) * ckw : nullptr;
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
            if (TRACE && ckc == 0) {                             // checkpoint = state after column j0-1 (a multiple of C)
                // (address from base + running offset: a pointer bumped after the stores waits for all of them to read it;
                //  streaming stores: a record is read back at most once, by another SM, much later)
                uint4 *const cq = reinterpret_cast<uint4 *>(ckp + ckoff);
#pragma unroll
                for (int g = 0; g < NR / 2; ++g) __stcs(cq + g, make_uint4(Mo[2 * g], Hl[2 * g], Mo[2 * g + 1], Hl[2 * g + 1]));
                ckoff += 2 * NR;
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
            ex[h] = (uint32_t)pci | ((uint32_t)pcj << 6) | ((uint32_t)pst << 18) | 0x80000000u;
            ey[h] = (uint32_t)sc;
        }
        *entp = make_uint4(ex[0], ey[0], ex[1], ey[1]);
    }
}

// ------------------------------------------------------------------------------------------------
// k_rounds<NR, EXACT, C>: pass 2 of the trace.  A task's traceback crosses 1.8 flag blocks on average but the longest of
// 64 crosses 3+, so a warp that took its 64 tasks through pass 2 in lockstep would spend a third of its time on a few
// lanes.  Instead a unit gets ONE round: every task's flag block holding its end cell is rebuilt from the checkpoint on
// its left and walked; the paths that end inside it are finished (statistics, result record), the others become ENTRIES
// of the warp's continuation queue in shared memory (5 words: which task, where the walk stands, the walk's
// accumulators).  A continuation round pops 64 entries -- from whatever units, the adapter is the same -- and does the
// same for each entry's next block; it runs whenever the queue holds 64, so every lane of a round works.  The queue is
// drained when the warp's adapter changes and at the end.
// dynamic smem: [W][NR*25] score tables, [W][NR] adapter codes, [W][5][PCF_QCAP] queues, then per thread (strided by
//               PCB_BLOCK) C*NR/4 flag words + 2*C/4 words of block bases.
// ------------------------------------------------------------------------------------------------
constexpr int QW = 5;
// queue words: 0 task = (unit - unit0) << 6 | slot        1 walk position = row | column << 6 | state << 18
//              2 end cell = row | column << 6 | (raw score + OFF) << 18
//              3 nD | matches << 6 | (kA + 1) << 12 | (kR + 1) << 24       4 jA | iR << 12      (WalkAcc, align_stats.h)
struct RoundCtx {
    int head, count;                               // queue state, warp-uniform
    int tbl_off, code_off, fl_off, q_off;          // word offsets into the CTA's dynamic shared memory (pointers through this
                                                   // struct would lose their address space: every LDS / STS a generic access)
    const uint32_t *ck;                            // checkpoints of the batch: [unit - unit0][lane][k][2 * NR]
    const uint4 *entries;
    uint64_t ckw;                                  // checkpoint words per thread and unit
    const PcbUnit *units;                          // explicit source: units of the batch (nullptr: cross product)
    const PcbTask *tasks;
    const int32_t *win_len;                        // cross product
    int64_t gbase;                                 // cross product: window group of batch unit u = u + gbase
    int32_t stride, n_ids_total, out_col;
    const uint8_t *arena;
    pcb_result *out;
    int32_t *err_flag;
    int n, r0, b, lane;
    uint32_t GEB, Bp;
};

template <int NR, bool EXACT, int C>
__device__ __noinline__ void trace_round(RoundCtx &q, const int64_t fresh_u /* >= 0: the entries of batch unit fresh_u; -1: pop */)
{
    constexpr int FW = NR / 4;
    // the context lives in local memory: take what the loops use into registers once
    const int lane = q.lane, n = q.n, r0 = q.r0, b = q.b;
    const uint64_t ckw = q.ckw;
    extern __shared__ uint32_t smem[];
    const uint32_t *const tbl = smem + q.tbl_off, *const s_code = smem + q.code_off;
    uint32_t *const s_fl = smem + q.fl_off, *const s_q = smem + q.q_off;
    const uint32_t GEB = q.GEB, Bp = q.Bp;
    const bool cross = q.units == nullptr;

    uint32_t tref[2], w2[2];
    int ci[2], cj[2], st[2], qpos[2];
    bool valid[2];
    if (fresh_u >= 0) {
        const uint4 e = q.entries[fresh_u * 32 + lane];
        tref[0] = ((uint32_t)fresh_u << 6) | (uint32_t)(2 * lane); tref[1] = tref[0] | 1u;
        valid[0] = (e.x >> 31) != 0u; valid[1] = (e.z >> 31) != 0u;
        ci[0] = (int)(e.x & 63u); cj[0] = (int)((e.x >> 6) & 0xFFFu); st[0] = (int)((e.x >> 18) & 7u);
        ci[1] = (int)(e.z & 63u); cj[1] = (int)((e.z >> 6) & 0xFFFu); st[1] = (int)((e.z >> 18) & 7u);
        w2[0] = (e.x & 0x3FFFFu) | (e.y << 18); w2[1] = (e.z & 0x3FFFFu) | (e.w << 18);
        qpos[0] = qpos[1] = -1;
    } else {
        const int take = min(q.count, 64);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            valid[h] = 2 * lane + h < take;
            qpos[h] = (q.head + 2 * lane + h) & (PCF_QCAP - 1);
            tref[h] = 0u; w2[h] = 0u; ci[h] = cj[h] = st[h] = 0;
            if (valid[h]) {
                tref[h] = s_q[qpos[h]];
                const uint32_t w1 = s_q[PCF_QCAP + qpos[h]];
                ci[h] = (int)(w1 & 63u); cj[h] = (int)((w1 >> 6) & 0xFFFu); st[h] = (int)((w1 >> 18) & 7u);
                w2[h] = s_q[2 * PCF_QCAP + qpos[h]];
            }
        }
        q.head = (q.head + take) & (PCF_QCAP - 1); q.count -= take;
    }
    // where the task's bases are
    uint64_t off[2]; int len[2]; int64_t widx[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        off[h] = 0; len[h] = 0; widx[h] = 0;
        if (!valid[h]) continue;
        if (cross) {
            const int64_t w = ((int64_t)(tref[h] >> 6) + q.gbase) * 64 + (int64_t)(tref[h] & 63u);
            widx[h] = w; off[h] = (uint64_t)w * (uint64_t)q.stride; len[h] = q.win_len[w];
        } else {
            const int64_t t = q.units[tref[h] >> 6].task0 + (int64_t)(tref[h] & 63u);
            widx[h] = t; off[h] = q.tasks[t].off; len[h] = q.tasks[t].len;
        }
    }
    WalkAcc acc[2];
    auto load_acc = [&]() {                        // the walk's accumulators are fetched where the walk starts: the column loop wants the registers
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            walk_acc_init(acc[h]);
            if (valid[h] && qpos[h] >= 0) {
                const uint32_t a = s_q[3 * PCF_QCAP + qpos[h]], c = s_q[4 * PCF_QCAP + qpos[h]];
                acc[h].nD = (int)(a & 63u); acc[h].matches = (int)((a >> 6) & 63u);
                acc[h].kA = (int)((a >> 12) & 0xFFFu) - 1; acc[h].kR = (int)(a >> 24) - 1;
                acc[h].jA = (int)(c & 0xFFFu); acc[h].iR = (int)(c >> 12);
            }
        }
    };

    // block of each half: columns k*C+1 .. k*C+C, holding its current column
    const bool go0 = valid[0] && ci[0] > 0 && cj[0] > 0, go1 = valid[1] && ci[1] > 0 && cj[1] > 0;
    const int k0 = go0 ? (cj[0] - 1) / C : 0, k1 = go1 ? (cj[1] - 1) / C : 0;
    const int jb0 = k0 * C, jb1 = k1 * C;
    const int tmax = __reduce_max_sync(0xffffffffu, max(go0 ? cj[0] - jb0 : 0, go1 ? cj[1] - jb1 : 0));
    if (tmax > 0) {
        const uint8_t *const seq0 = q.arena + off[0], *const seq1 = q.arena + off[1];
        const int lim0 = max(len[0] - 1, 0), lim1 = max(len[1] - 1, 0);
        // the block's read bases (C per half), then the checkpointed column the block starts from
        uint32_t q0[C / 4], q1[C / 4];
#pragma unroll
        for (int g = 0; g < C / 4; ++g) {
            q0[g] = join4(load4(seq0, jb0 + 4 * g, lim0)); q1[g] = join4(load4(seq1, jb1 + 4 * g, lim1));
            s_fl[(C * FW + g) * PCB_BLOCK] = q0[g]; s_fl[(C * FW + C / 4 + g) * PCB_BLOCK] = q1[g];   // for the walk
        }
        uint32_t Mo[NR], Hl[NR];
        {
            // a task's checkpoints were written by the thread that ran its pass 1 (task >> 1 = batch unit * 32 + lane), two
            // tasks per word; a record is 2*NR contiguous words, so any lane can fetch any task's with full sectors
            const uint4 *p0 = reinterpret_cast<const uint4 *>(q.ck + ((uint64_t)(tref[0] >> 1) * ckw + (uint64_t)k0 * 2ull * NR));
            const uint4 *p1 = reinterpret_cast<const uint4 *>(q.ck + ((uint64_t)(tref[1] >> 1) * ckw + (uint64_t)k1 * 2ull * NR));
            const uint32_t sel = ((tref[0] & 1u) ? 0x0032u : 0x0010u) | ((tref[1] & 1u) ? 0x7600u : 0x5400u);
#pragma unroll
            for (int g = 0; g < NR / 2; ++g) {
                const uint4 a = __ldcs(p0 + g), c = __ldcs(p1 + g);
                Mo[2 * g] = prmt(a.x, c.x, sel); Hl[2 * g] = prmt(a.y, c.y, sel);
                Mo[2 * g + 1] = prmt(a.z, c.z, sel); Hl[2 * g + 1] = prmt(a.w, c.w, sel);
            }
        }
        uint32_t rT = pk(B0 + b * jb0, B0 + b * jb1);
        const bool pend0 = go0 && (st[0] & 4), pend1 = go1 && (st[1] & 4);
        const bool anypend = __any_sync(0xffffffffu, pend0 || pend1);
        int s0 = st[0] & 3, s1 = st[1] & 3;
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
            col_flags<NR, EXACT>(Mo, Hl, tp, rowTp, rT, GEB, s_fl + (t - 1) * FW * PCB_BLOCK, Ml, Vl, Hlast);
            if (anypend) {
                // end cell in the last adapter row: V == M / H == M of that cell decide how the trace starts
                // (_correctTraceValue, dp_algorithm_impl.h:1169-1185)
                const uint32_t xv = Ml ^ Vl, xh = Ml ^ Hlast;
                if (pend0 && jb0 + t == cj[0]) s0 = ((xv & 0xFFFFu) == 0) ? 1 : (((xh & 0xFFFFu) == 0) ? 2 : 0);
                if (pend1 && jb1 + t == cj[1]) s1 = ((xv >> 16) == 0) ? 1 : (((xh >> 16) == 0) ? 2 : 0);
            }
        }
        __syncwarp();
        load_acc();
        // walk (dp_traceback_impl.h:379-451, :498-545) inside the block: both halves together, one path op per half and trip
        {
            int i0 = ci[0], j0 = cj[0], i1 = ci[1], j1 = cj[1];
            WalkAcc a0 = acc[0], a1 = acc[1];
            bool act0 = go0 && j0 > jb0, act1 = go1 && j1 > jb1;
            while (act0 || act1) {
                if (act0) { walk_step<FW, C>(s_fl, s_code, 0, jb0, r0, i0, j0, s0, a0); act0 = i0 > 0 && j0 > jb0; }
                if (act1) { walk_step<FW, C>(s_fl, s_code, 1, jb1, r0, i1, j1, s1, a1); act1 = i1 > 0 && j1 > jb1; }
            }
            ci[0] = i0; cj[0] = j0; acc[0] = a0; ci[1] = i1; cj[1] = j1; acc[1] = a1;
        }
        st[0] = s0; st[1] = s1;                                  // the pending bit is spent
    } else {
        load_acc();
    }
    __syncwarp();                                                // every lane has read its entry before a queue slot is written again

    // paths that reached row 0 / column 0 are complete (this also covers the empty path); the others go (back) into the queue
    bool keep[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        keep[h] = valid[h] && ci[h] > 0 && cj[h] > 0;
        if (!valid[h] || keep[h]) continue;
        int col0 = 0, mtrue = len[h];
        int64_t oidx;
        if (cross) oidx = widx[h] * q.n_ids_total + q.out_col;
        else { const PcbTask &k = q.tasks[widx[h]]; col0 = k.col0; mtrue = k.m_true; oidx = k.out_idx; }
        const int endi = (int)(w2[h] & 63u), endj = (int)((w2[h] >> 6) & 0xFFFu);
        if (cj[h] == 0 && ci[h] > 0 && col0 > 0) atomicOr(q.err_flag, 1);      // window too narrow for this path
        PathAcc pa; walk_acc_finish(acc[h], endi, endj, ci[h], cj[h], pa);
        PathStats ps;
        path_finalize(pa, (int64_t)col0 + cj[h], ci[h], (int64_t)mtrue - (col0 + endj), n - endi, mtrue, n, ps);
        pcb_result res;
        res.read_start = ps.read_start; res.read_end = ps.read_end;
        res.adapter_start = ps.adapter_start; res.adapter_end = ps.adapter_end;
        res.raw_score = (int)(w2[h] >> 18) - OFF; res.matches = ps.matches;
        res.aligned_len = ps.aligned_len; res.full_len = ps.full_len;
        q.out[oidx] = res;
    }
    const uint32_t m0 = __ballot_sync(0xffffffffu, keep[0]), m1 = __ballot_sync(0xffffffffu, keep[1]);
    const uint32_t lt = (1u << lane) - 1u;
    const int tail = q.head + q.count;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        if (!keep[h]) continue;
        const int pos = (tail + (h ? __popc(m0) + __popc(m1 & lt) : __popc(m0 & lt))) & (PCF_QCAP - 1);
        s_q[pos] = tref[h];
        s_q[PCF_QCAP + pos] = (uint32_t)ci[h] | ((uint32_t)cj[h] << 6) | ((uint32_t)st[h] << 18);
        s_q[2 * PCF_QCAP + pos] = w2[h];
        s_q[3 * PCF_QCAP + pos] = (uint32_t)acc[h].nD | ((uint32_t)acc[h].matches << 6) | ((uint32_t)(acc[h].kA + 1) << 12) | ((uint32_t)(acc[h].kR + 1) << 24);
        s_q[4 * PCF_QCAP + pos] = (uint32_t)acc[h].jA | ((uint32_t)acc[h].iR << 12);
    }
    q.count += __popc(m0) + __popc(m1);
    __syncwarp();
}

template <int NR, bool EXACT, int C>
__global__ void __launch_bounds__(PCB_BLOCK, PCF_MINB_P2(NR)) k_rounds(const PcbLaunch L)
{
    extern __shared__ uint32_t smem[];
    constexpr int FW = NR / 4;
    constexpr int W = PCB_BLOCK / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *const tbl = smem + warp * (NR * TP);
    uint32_t *const s_code = smem + W * (NR * TP) + warp * NR;

    const int b = -L.sc.gap_open;
    const int s_match = L.sc.match + 2 * b, s_mis = L.sc.mismatch + 2 * b;
    const int64_t nwarps = (int64_t)gridDim.x * W, gw = (int64_t)blockIdx.x * W + warp;
    const bool cross = L.src.units == nullptr;
    const int64_t groups = cross ? (L.src.n_windows + 63) / 64 : 0;
    const int64_t nu = L.unit1 - L.unit0;
    int64_t u_begin, u_end, u_step;
    if (cross) { u_begin = L.unit0 + gw * nu / nwarps; u_end = L.unit0 + (gw + 1) * nu / nwarps; u_step = 1; }
    else       { u_begin = L.unit0 + gw; u_end = L.unit1; u_step = nwarps; }

    RoundCtx q;
    q.head = 0; q.count = 0;
    q.tbl_off = (int)(tbl - smem); q.code_off = (int)(s_code - smem);
    q.q_off = W * (NR * TP + NR) + warp * (QW * PCF_QCAP);
    q.fl_off = W * (NR * TP + NR + QW * PCF_QCAP) + (int)threadIdx.x;
    q.ck = L.scratch; q.entries = L.entries; q.ckw = ckw_of(L.mmax, C, NR);
    q.units = cross ? nullptr : L.src.units + L.unit0; q.tasks = L.src.tasks;
    q.win_len = L.src.win_len; q.gbase = 0; q.stride = L.src.stride; q.n_ids_total = L.src.n_ids_total; q.out_col = 0;
    q.arena = L.arena; q.out = L.out; q.err_flag = L.err_flag;
    q.n = 0; q.r0 = 0; q.b = b; q.lane = lane;
    q.GEB = pk2(L.sc.gap_extend + b); q.Bp = pk2(b);
    (void)FW;

    int32_t tbl_adapter = -1;
    int64_t cur_k = -1;
    for (int64_t unit = u_begin; unit < u_end; unit += u_step) {
        int32_t adapter;
        int64_t k = 0;
        if (cross) { k = unit / groups; adapter = L.src.ids[k]; }
        else adapter = L.src.units[unit].adapter;
        const int64_t aoff = L.ad.off[adapter];
        const int n = (int)(L.ad.off[adapter + 1] - aoff);
        if (n <= 0 || n > NR || (EXACT && n != NR)) continue;          // pass 1 wrote the failed records and raised the flag
        const int r0 = EXACT ? 0 : NR - n;
        if (adapter != tbl_adapter || k != cur_k) {
            while (q.count > 0) trace_round<NR, EXACT, C>(q, -1);     // the queue's entries belong to the old adapter / result column
            if (cross) { q.gbase = L.unit0 - k * groups; q.out_col = L.src.id_pos ? L.src.id_pos[k] : (int32_t)k; }
            cur_k = k;
        }
        if (adapter != tbl_adapter) {
            q.n = n; q.r0 = r0;
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
        trace_round<NR, EXACT, C>(q, unit - L.unit0);
        while (q.count >= 64) trace_round<NR, EXACT, C>(q, -1);
    }
    while (q.count > 0) trace_round<NR, EXACT, C>(q, -1);
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// host-side launchers, by kernel class
// ------------------------------------------------------------------------------------------------
static size_t fast_smem(int kind, int nr, int c)
{
    const size_t W = PCB_BLOCK / 32;
    if (kind == 2)       // rounds: tables, codes, queues, then flags + block bases per thread
        return (W * (size_t)(nr * TP + nr + QW * PCF_QCAP) + (size_t)PCB_BLOCK * ((size_t)c * (nr / 4) + 2 * (c / 4))) * sizeof(uint32_t);
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
    default: { constexpr int NR = 52; constexpr bool EX = false; constexpr int C = PCF_C_LARGE; __VA_ARGS__; } break; \
    }

cudaError_t pcf_launch_scan(const PcbLaunch &L, int cls, cudaStream_t st)
{
    if (L.unit1 <= L.unit0) return cudaSuccess;
    if (cls < 0 || cls >= PCF_NCLS) return cudaErrorInvalidValue;
    PCF_BY_CLASS(return (launch_fast<0, NR, EX, C>(L, L.grid, st)))
    return cudaSuccess;
}

cudaError_t pcf_launch_trace(const PcbLaunch &L, int cls, int grid_rounds, cudaStream_t st)
{
    if (L.unit1 <= L.unit0) return cudaSuccess;
    if (cls < 0 || cls >= PCF_NCLS) return cudaErrorInvalidValue;
    cudaError_t e = cudaSuccess;
    PCF_BY_CLASS(e = launch_fast<1, NR, EX, C>(L, L.grid, st); if (e == cudaSuccess) e = launch_rounds<NR, EX, C>(L, grid_rounds, st))
    return e;
}

int pcf_occupancy_grid(int kind, int cls)
{
    int dev = 0, sms = 0, per = 1;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (kind == 0)      { PCF_BY_CLASS(per = (occ_one<0, NR, EX, C>())) }
    else if (kind == 1) { PCF_BY_CLASS(per = (occ_one<1, NR, EX, C>())) }
    else                { PCF_BY_CLASS(per = (occ_one<2, NR, EX, C>())) }
    return sms * per;
}
