/*
 * align_stats.h -- ScoredAlignment statistics from an alignment path, without materialising the
 * two gapped rows.  Shared by the CUDA traceback (device) and pcb_stats_from_path (host, tests).
 *
 * What it reproduces: porechop_abi/src/alignment.cpp:5-110 of the reference.  The reference prints
 * both rows to strings and scans them five times; here the same numbers are derived in closed form
 * from the layout the traceback produces (seqan/align/dp_traceback_impl.h:498-545 and
 * dp_traceback_adaptor.h:60-118):
 *
 *     [head_h read-only cols][head_v adapter-only cols][core path, K cols][tail_h][tail_v]
 *
 * Every column holds at least one base, matches only occur in 'D' columns, and every 'D' column lies
 * inside both identity ranges, so alignedMatchCount == fullAdapterMatchCount == #matching D columns.
 */
#ifndef PCB_ALIGN_STATS_H
#define PCB_ALIGN_STATS_H

#include <stdint.h>

#if defined(__CUDACC__)
#define PCB_HD __host__ __device__ __forceinline__
#else
#define PCB_HD static inline
#endif

/* Accumulated while walking the core path BACK TO FRONT (k_rev = 0 is the last column of the core). */
struct PathAcc {
    int32_t K;               /* core columns so far                                   */
    int32_t matches;         /* D columns with equal bases                             */
    int32_t kA_last_rev;     /* first (in reverse) core column with an adapter base    */
    int32_t kA_first_rev;    /* last  (in reverse) core column with an adapter base    */
    int32_t kR_last_rev;
    int32_t kR_first_rev;
    int32_t A_last_isD;      /* is that column a D (i.e. does it also hold a read base) */
    int32_t R_last_isD;
    int32_t A_last_match;    /* ... and do its two bases match (general kernels only; the compatibility flag needs it) */
    int32_t R_last_match;
};

PCB_HD void path_acc_init(PathAcc &a)
{
    a.K = 0; a.matches = 0;
    a.kA_last_rev = -1; a.kA_first_rev = -1; a.kR_last_rev = -1; a.kR_first_rev = -1;
    a.A_last_isD = 0; a.R_last_isD = 0; a.A_last_match = 0; a.R_last_match = 0;
}

/* op: 0 = D, 1 = V (adapter base, read gap), 2 = H (read base, adapter gap) */
PCB_HD void path_acc_step(PathAcc &a, int op, int is_match)
{
    const int k = a.K;
    if (op != 2) { if (a.kA_last_rev < 0) { a.kA_last_rev = k; a.A_last_isD = (op == 0); a.A_last_match = (op == 0 && is_match); } a.kA_first_rev = k; }
    if (op != 1) { if (a.kR_last_rev < 0) { a.kR_last_rev = k; a.R_last_isD = (op == 0); a.R_last_match = (op == 0 && is_match); } a.kR_first_rev = k; }
    if (op == 0 && is_match) a.matches++;
    a.K = k + 1;
}

/* Accumulation of the fast kernels' walk (pcb_fast.cu), packed so that a step has no branch; the step indices PathAcc
 * wants are recovered at the end (rows and columns only ever decrease along the reverse walk):
 *   kAR : low half = 1 + max over the ops that hold an adapter base (D, V) of 2 * column + isV, column = the one before the
 *         op (the FIRST such op has the largest; 0 = none yet); high half = the same for read bases (D, H) with rows
 *   tHV : low = H ops since the last op with an adapter base, high = V ops since the last op with a read base
 *   nHV : low = H ops, high = V ops;   op : gap openings
 * pos2 = (2 * column + 1) | (2 * row + 1) << 16 of the cell the op leaves.  isD / isV / isH: exactly one is 1. */
struct WalkAcc { uint32_t kAR, tHV, nHV; int32_t op; };
PCB_HD void walk_acc_init(WalkAcc &a) { a.kAR = 0; a.tHV = 0; a.nHV = 0; a.op = 0; }
PCB_HD uint32_t walk_max16x2(uint32_t a, uint32_t b)
{
#if defined(__CUDA_ARCH__)
    return __vmaxs2(a, b);
#else
    const uint32_t lo = (a & 0xFFFFu) > (b & 0xFFFFu) ? (a & 0xFFFFu) : (b & 0xFFFFu), hi = (a >> 16) > (b >> 16) ? (a >> 16) : (b >> 16);
    return lo | (hi << 16);
#endif
}
PCB_HD void walk_acc_step(WalkAcc &a, uint32_t &pos2, uint32_t isD, uint32_t isV, uint32_t isH, uint32_t ext)
{
    const uint32_t inc = isH | (isV << 16), incs = isV | (isH << 16);
    const uint32_t keep = inc * 0xFFFFu;                          /* H: 0x0000FFFF, V: 0xFFFF0000, D: 0 */
    a.kAR = walk_max16x2(a.kAR, (pos2 + incs) & ~keep);
    a.tHV = (a.tHV & keep) + inc;
    a.nHV += inc;
    pos2 = pos2 - 0x00020002u + 2u * incs;
    a.op += (int32_t)((isD | ext) ^ 1u);
}
/* (endi, endj): cell the walk started from; (ci, cj): cell it stopped at.  matches: counted by the caller, or derived from
 * the end cell's score (every D op scores match or mismatch, every gap of g ops gapOpen + (g - 1) gapExtend). */
PCB_HD void walk_acc_finish(const WalkAcc &w, int endi, int endj, int ci, int cj, int matches, PathAcc &a)
{
    const int nH = (int)(w.nHV & 0xFFFFu), nV = (int)(w.nHV >> 16), kA = (int)(w.kAR & 0xFFFFu) - 1, kR = (int)(w.kAR >> 16) - 1;
    const int nD = (endi - ci) - nV;
    a.K = nD + nV + nH; a.matches = matches;
    a.kA_last_rev = -1; a.kA_first_rev = -1; a.kR_last_rev = -1; a.kR_first_rev = -1; a.A_last_isD = 0; a.R_last_isD = 0;
    a.A_last_match = 0; a.R_last_match = 0;
    if (kA >= 0) { a.kA_last_rev = endj - (kA >> 1); a.A_last_isD = !(kA & 1); a.kA_first_rev = a.K - 1 - (int)(w.tHV & 0xFFFFu); }
    if (kR >= 0) { a.kR_last_rev = endi - (kR >> 1); a.R_last_isD = !(kR & 1); a.kR_first_rev = a.K - 1 - (int)(w.tHV >> 16); }
    (void)cj;
}
PCB_HD int walk_acc_matches(const WalkAcc &w, int endi, int endj, int ci, int cj, int score, int match, int mismatch,
                            int gap_open, int gap_extend, int *ok)
{
    const int nH = (int)(w.nHV & 0xFFFFu), nV = (int)(w.nHV >> 16), nD = (endi - ci) - nV;
    const int num = score - mismatch * nD - gap_open * w.op - gap_extend * (nV + nH - w.op), den = match - mismatch;
    *ok = nD == (endj - cj) - nH && num >= 0 && den > 0 && num % den == 0;
    return den > 0 ? num / den : 0;
}

struct PathStats {
    int32_t read_start, read_end, adapter_start, adapter_end, matches, aligned_len, full_len;
    int32_t end_match;       /* the last column of the aligned region holds two equal bases (general kernels only) */
};

/* m, n: full read / adapter lengths (head_h + #D + #H + tail_h == m, head_v + #D + #V + tail_v == n). */
PCB_HD void path_finalize(const PathAcc &a, int64_t head_h, int64_t head_v, int64_t tail_h, int64_t tail_v,
                          int64_t m, int64_t n, PathStats &o)
{
    const int64_t K = a.K;
    const int64_t c0 = head_h + head_v, c1 = c0 + K, L = c1 + tail_h + tail_v;
    const bool hasA = a.kA_last_rev >= 0, hasR = a.kR_last_rev >= 0;
    /* first / last column holding a read base, an adapter base (alignment.cpp:26-70) */
    const int64_t firstRead = head_h > 0 ? 0 : (hasR ? c0 + (K - 1 - a.kR_first_rev) : c1);
    const int64_t firstAd   = head_v > 0 ? head_h : (hasA ? c0 + (K - 1 - a.kA_first_rev) : c1 + tail_h);
    const int64_t lastRead  = tail_h > 0 ? c1 + tail_h - 1 : (hasR ? c0 + (K - 1 - a.kR_last_rev) : head_h - 1);
    const int64_t lastAd    = tail_v > 0 ? L - 1 : (hasA ? c0 + (K - 1 - a.kA_last_rev) : c0 - 1);
    const int64_t startPos = firstRead > firstAd ? firstRead : firstAd;
    const int64_t endPos   = lastRead < lastAd ? lastRead : lastAd;
    /* bases before startPos (alignment.cpp:95-98): columns before the later of the two "first" columns
     * hold only the other sequence's bases */
    o.read_start    = (int32_t)(firstAd > firstRead ? firstAd - firstRead : 0);
    o.adapter_start = (int32_t)(firstRead > firstAd ? firstRead - firstAd : 0);
    /* bases before endPos (alignment.cpp:99-102) = total - bases at or after endPos */
    int64_t readAtOrAfter, adAtOrAfter;
    if (lastRead <= lastAd) {
        readAtOrAfter = 1;
        const bool colHasAd = (tail_h > 0) ? false : (hasR ? (a.R_last_isD != 0) : false);
        adAtOrAfter = (lastAd - lastRead) + (colHasAd ? 1 : 0);
    } else {
        adAtOrAfter = 1;
        const bool colHasRead = (tail_v > 0) ? false : (hasA ? (a.A_last_isD != 0) : false);
        readAtOrAfter = (lastRead - lastAd) + (colHasRead ? 1 : 0);
    }
    o.read_end    = (int32_t)(m - readAtOrAfter);
    o.adapter_end = (int32_t)(n - adAtOrAfter);
    o.matches     = a.matches;
    /* the column endPos: the last read base's column or the last adapter base's, whichever comes first */
    o.end_match   = (lastRead <= lastAd) ? ((tail_h > 0 || !hasR) ? 0 : a.R_last_match) : ((tail_v > 0 || !hasA) ? 0 : a.A_last_match);
    o.aligned_len = (int32_t)(endPos - startPos + 1);
    o.full_len    = (int32_t)(lastAd - firstAd + 1);
}

#endif
