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
};

PCB_HD void path_acc_init(PathAcc &a)
{
    a.K = 0; a.matches = 0;
    a.kA_last_rev = -1; a.kA_first_rev = -1; a.kR_last_rev = -1; a.kR_first_rev = -1;
    a.A_last_isD = 0; a.R_last_isD = 0;
}

/* op: 0 = D, 1 = V (adapter base, read gap), 2 = H (read base, adapter gap) */
PCB_HD void path_acc_step(PathAcc &a, int op, int is_match)
{
    const int k = a.K;
    if (op != 2) { if (a.kA_last_rev < 0) { a.kA_last_rev = k; a.A_last_isD = (op == 0); } a.kA_first_rev = k; }
    if (op != 1) { if (a.kR_last_rev < 0) { a.kR_last_rev = k; a.R_last_isD = (op == 0); } a.kR_first_rev = k; }
    if (op == 0 && is_match) a.matches++;
    a.K = k + 1;
}

/* Cheaper accumulation for the fast kernels' walk: instead of step indices it keeps cell coordinates, and the step
 * indices PathAcc wants are recovered at the end (rows and columns only ever decrease along the reverse walk):
 *   kA = max over ops holding an adapter base (D, V) of 2*j + isV, j = column before the op: the FIRST such op has the
 *        largest j, and if V ops share that column they come before the D that leaves it;
 *   kR = the same for read bases (D, H) with rows: 2*i + isH;
 *   jA / iR = column / row after the LAST op holding an adapter / read base.  */
struct WalkAcc {
    int32_t nD, matches, kA, kR, jA, iR;
};
PCB_HD void walk_acc_init(WalkAcc &a) { a.nD = 0; a.matches = 0; a.kA = -1; a.kR = -1; a.jA = 0; a.iR = 0; }
PCB_HD void walk_acc_D(WalkAcc &a, int i, int j, int is_match)
{
    a.nD++; a.matches += is_match;
    a.kA = a.kA > 2 * j ? a.kA : 2 * j; a.kR = a.kR > 2 * i ? a.kR : 2 * i;
    a.jA = j - 1; a.iR = i - 1;
}
PCB_HD void walk_acc_V(WalkAcc &a, int i, int j) { (void)i; a.kA = a.kA > 2 * j + 1 ? a.kA : 2 * j + 1; a.jA = j; }
PCB_HD void walk_acc_H(WalkAcc &a, int i, int j) { (void)j; a.kR = a.kR > 2 * i + 1 ? a.kR : 2 * i + 1; a.iR = i; }
/* (endi, endj): cell the walk started from; (ci, cj): cell it stopped at. */
PCB_HD void walk_acc_finish(const WalkAcc &w, int endi, int endj, int ci, int cj, PathAcc &a)
{
    const int K = (endi - ci) + (endj - cj) - w.nD;
    a.K = K; a.matches = w.matches;
    a.kA_last_rev = -1; a.kA_first_rev = -1; a.kR_last_rev = -1; a.kR_first_rev = -1; a.A_last_isD = 0; a.R_last_isD = 0;
    if (w.kA >= 0) { a.kA_last_rev = endj - (w.kA >> 1); a.A_last_isD = !(w.kA & 1); a.kA_first_rev = K - 1 - (w.jA - cj); }
    if (w.kR >= 0) { a.kR_last_rev = endi - (w.kR >> 1); a.R_last_isD = !(w.kR & 1); a.kR_first_rev = K - 1 - (w.iR - ci); }
}

struct PathStats {
    int32_t read_start, read_end, adapter_start, adapter_end, matches, aligned_len, full_len;
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
    o.aligned_len = (int32_t)(endPos - startPos + 1);
    o.full_len    = (int32_t)(lastAd - firstAd + 1);
}

#endif
