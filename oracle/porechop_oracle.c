/*
 * porechop_oracle.c -- CPU oracle (TEST INFRASTRUCTURE ONLY, see porechop_oracle.h).
 *
 * A literal, unoptimised restatement of the reference's single-pair alignment.  Every block
 * cites the reference lines it follows; paths are relative to /root/reference/porechop_abi/ and
 * "seqan/" means include/seqan/ (vendored SeqAn 2.4.0).  No code was copied: SeqAn's templates
 * are collapsed to the one instantiation the reference uses
 *   globalAlignment(Align<Dna5String,ArrayGaps>, Score<int,Simple>, AlignConfig<true,true,true,true>)
 *   (src/adapter_align.cpp:20-27)
 * i.e. FreeEndGaps_<True,True,True,True>, BandOff, TracebackOn<SingleTrace,GapsLeft>, with
 * AffineGaps (Gotoh) when gapOpen != gapExtend and LinearGaps (NeedlemanWunsch) when equal
 * (seqan/align/global_alignment_unbanded.h:210-221).
 */
#include "porechop_oracle.h"

#include <limits.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* seqan/align/dp_profile.h:120-133 (TraceValue_ bit values) */
enum {
    TB_NONE = 0, TB_DIAG = 1, TB_H = 2, TB_V = 4, TB_H_OPEN = 8, TB_V_OPEN = 16,
    TB_MAX_FROM_H = 32, TB_MAX_FROM_V = 64
};

/* seqan/basic/alphabet_residue_tabs.h:113-140: A/a=0 C/c=1 G/g=2 T/t/U/u=3, every other byte = 4 (N) */
static unsigned char dna5_of(unsigned char c)
{
    switch (c) {
    case 'A': case 'a': return 0;
    case 'C': case 'c': return 1;
    case 'G': case 'g': return 2;
    case 'T': case 't': case 'U': case 'u': return 3;
    default: return 4;
    }
}

/* seqan/align/dp_cell.h:144-145: "infinity" is INT_MIN / 2 so that adding gap scores cannot wrap */
#define DP_NEG_INF (INT_MIN / 2)

static void failed_result(pc_oracle_result *out)
{
    /* ScoredAlignment ctor initialises only these (src/alignment.cpp:7-8); the reference then
     * returns early (:20-21).  Other members are uninitialised there; the values below are what
     * the reference build prints in practice for empty input (see tests/golden). */
    out->read_start = -1; out->read_end = 0; out->adapter_start = -1; out->adapter_end = 0;
    out->raw_score = INT_MIN; out->aligned_matches = 0; out->aligned_len = 0;
    out->full_matches = 0; out->full_len = 0;
}

/* path_ops / path_flags / path_meta may be NULL; when given they receive the core path front to back
 * ('D','V','H'), a 1 per column whose two bases are equal, and {head_h, head_v, tail_h, tail_v, score}. */
static int align_impl(const char *read_seq, long m, const char *adapter_seq, long n,
                      int match, int mismatch, int gap_open, int gap_extend,
                      pc_oracle_result *out, char *path_ops, char *path_flags, long *path_meta, long *path_n)
{
    if (path_n) *path_n = 0;
    if (m <= 0 || n <= 0) { failed_result(out); return 0; }

    const int affine = (gap_open != gap_extend);   /* global_alignment_unbanded.h:216-219 */
    const long rows = n + 1;

    unsigned char *h = (unsigned char *)malloc((size_t)m);
    unsigned char *v = (unsigned char *)malloc((size_t)n);
    int *colM = (int *)malloc(sizeof(int) * (size_t)rows);    /* sparse score matrix: one column */
    int *colH = (int *)malloc(sizeof(int) * (size_t)rows);    /*  (dp_algorithm_impl.h:1380-1404) */
    unsigned char *trace = (unsigned char *)malloc((size_t)rows * (size_t)(m + 1)); /* full trace matrix */
    char *ops = (char *)malloc((size_t)(m + n + 2));
    if (!h || !v || !colM || !colH || !trace || !ops) {
        free(h); free(v); free(colM); free(colH); free(trace); free(ops);
        return -1;
    }
    for (long j = 0; j < m; ++j) h[j] = dna5_of((unsigned char)read_seq[j]);      /* adapter_align.cpp:15 */
    for (long i = 0; i < n; ++i) v[i] = dna5_of((unsigned char)adapter_seq[i]);   /* adapter_align.cpp:16 */

    /* Scout: best tracked cell, strict '>' so the first one in visiting order wins
     * (seqan/align/dp_scout.h:167-180).  Tracked cells = last cell of every column (free last row)
     * and every cell of the final column (free last column)  (dp_meta_info.h:205-216). */
    int best = DP_NEG_INF, bestH = DP_NEG_INF, bestV = DP_NEG_INF;
    long bestI = 0, bestJ = 0;

    /* ---- initial column: every cell RecursionDirectionZero (dp_meta_info.h:186-195, dp_formula.h:275-289) */
    for (long i = 0; i <= n; ++i) { colM[i] = 0; colH[i] = DP_NEG_INF; trace[i] = TB_NONE; }
    if (colM[n] > best) { best = colM[n]; bestH = colH[n]; bestV = DP_NEG_INF; bestI = n; bestJ = 0; }

    /* ---- inner + final columns, column-wise, top to bottom (dp_algorithm_impl.h:488-574, :344-435) */
    for (long j = 1; j <= m; ++j) {
        unsigned char *tcol = trace + (size_t)j * (size_t)rows;
        const int lastcol = (j == m);
        /* first cell: Zero (free first row); caches are fresh per column (dp_algorithm_impl.h:449-454) */
        int diag = colM[0];           /* M[i-1][j-1] */
        int upM = 0;                  /* M[i-1][j]   */
        int upV = DP_NEG_INF;         /* V[i-1][j]   */
        colM[0] = 0; tcol[0] = TB_NONE;        /* colH[0] stays at -inf: Zero never writes it */
        if (lastcol && colM[0] > best) { best = 0; bestH = colH[0]; bestV = upV; bestI = 0; bestJ = j; }

        for (long i = 1; i <= n; ++i) {
            const int leftM = colM[i], leftH = colH[i];
            const int sub = (h[j - 1] == v[i - 1]) ? match : mismatch;   /* score_simple.h: N==N is a match */
            int cur, curH = DP_NEG_INF, curV = DP_NEG_INF;
            unsigned char tv;
            if (affine) {
                /* dp_formula_affine.h:65-124; every _maxScore keeps the LEFT operand on ties (dp_formula.h:147-163) */
                const int hExt = leftH + gap_extend, hOpen = leftM + gap_open;
                if (hExt < hOpen) { curH = hOpen; tv = TB_H_OPEN; } else { curH = hExt; tv = TB_H; }
                const int vExt = upV + gap_extend, vOpen = upM + gap_open;
                if (vExt < vOpen) { curV = vOpen; tv |= TB_V_OPEN; } else { curV = vExt; tv |= TB_V; }
                unsigned char tv2; int gapBest;
                if (curV < curH) { gapBest = curH; tv2 = TB_MAX_FROM_H; } else { gapBest = curV; tv2 = TB_MAX_FROM_V; }
                const int d = diag + sub;
                if (d < gapBest) { cur = gapBest; tv = (unsigned char)(tv2 | tv); }
                else             { cur = d;       tv = (unsigned char)(TB_DIAG | tv); }
            } else {
                /* dp_formula_linear.h:64-103 */
                const int vS = upM + gap_extend, hS = leftM + gap_extend;
                int gapBest;
                if (vS < hS) { gapBest = hS; tv = TB_H | TB_MAX_FROM_H; } else { gapBest = vS; tv = TB_V | TB_MAX_FROM_V; }
                const int d = diag + sub;
                if (d < gapBest) { cur = gapBest; } else { cur = d; tv = TB_DIAG; }
            }
            tcol[i] = tv;
            diag = leftM;
            colM[i] = cur; colH[i] = curH;
            upM = cur; upV = curV;
            if ((i == n || lastcol) && cur > best) {     /* _computeCell tracking, dp_algorithm_impl.h:302-314 */
                best = cur; bestH = curH; bestV = curV; bestI = i; bestJ = j;
            }
        }
    }

    /* ---- _finishAlignment -> _correctTraceValue, affine only (dp_algorithm_impl.h:1169-1185, :1320-1334) */
    unsigned char *cell = trace + (size_t)bestJ * (size_t)rows + bestI;
    if (affine) {
        if (bestV == best)      { *cell = (unsigned char)((*cell & ~TB_DIAG) | TB_MAX_FROM_V); }
        else if (bestH == best) { *cell = (unsigned char)((*cell & ~TB_DIAG) | TB_MAX_FROM_H); }
    }

    /* ---- traceback (dp_traceback_impl.h:498-545, :379-451).  We collect one op per alignment
     *      column, back to front: 'D' diagonal, 'V' adapter base over a read gap, 'H' read base over
     *      an adapter gap.  Segment merging in the reference is bookkeeping only. */
    long nops = 0;
    long ci = bestI, cj = bestJ;
    /* trace tail (TraceTail_: free end gaps), recorded first = rightmost in the alignment (:513-521) */
    long tailV = n - ci, tailH = m - cj;
    unsigned char tv = *cell;
    if (affine) {    /* _retrieveInitialTraceDirection with PreferGapsAtEnd_ (:458-471) */
        if (tv & TB_MAX_FROM_V)      tv &= (TB_V | TB_V_OPEN | TB_MAX_FROM_V);
        else if (tv & TB_MAX_FROM_H) tv &= (TB_H | TB_H_OPEN | TB_MAX_FROM_H);
    }
#define TRACE_AT(I, J) trace[(size_t)(J) * (size_t)rows + (size_t)(I)]
    while (!(cj <= 0 || ci <= 0) && tv != TB_NONE) {            /* _hasReachedEnd, :131-135 */
        if (tv & TB_DIAG) {                                      /* GapsLeft order of tests, :389-393 */
            --ci; --cj; ops[nops++] = 'D'; tv = TRACE_AT(ci, cj);
        } else if ((tv & TB_MAX_FROM_V) && (tv & TB_V)) {        /* _doTracebackGoVertical, :211-249 */
            if (affine) {
                while ((!(tv & TB_V_OPEN) || (tv & TB_V)) && ci != 1) {
                    --ci; ops[nops++] = 'V'; tv = TRACE_AT(ci, cj);
                }
            }
            --ci; ops[nops++] = 'V'; tv = TRACE_AT(ci, cj);
        } else if ((tv & TB_MAX_FROM_V) && (tv & TB_V_OPEN)) {   /* _doTracebackMaxFromVertical, :255-277 */
            --ci; ops[nops++] = 'V'; tv = TRACE_AT(ci, cj);
        } else if ((tv & TB_MAX_FROM_H) && (tv & TB_H)) {        /* _doTracebackGoHorizontal, :283-321 */
            if (affine) {
                while ((!(tv & TB_H_OPEN) || (tv & TB_H)) && cj != 1) {
                    --cj; ops[nops++] = 'H'; tv = TRACE_AT(ci, cj);
                }
            }
            --cj; ops[nops++] = 'H'; tv = TRACE_AT(ci, cj);
        } else if ((tv & TB_MAX_FROM_H) && (tv & TB_H_OPEN)) {   /* _doTracebackMaxFromHorizontal, :327-349 */
            --cj; ops[nops++] = 'H'; tv = TRACE_AT(ci, cj);
        } else {
            break;   /* NONE or undefined: the reference returns / asserts (:412-419) */
        }
    }
#undef TRACE_AT
    const long headV = ci, headH = cj;   /* TraceHead_: leading free gaps (:529-536) */

    /* ---- gapped rows as _adaptTraceSegmentsTo builds them (dp_traceback_adaptor.h:60-118): segments are
     *      consumed from the last recorded to the first, so the layout left to right is
     *      [head H][head V][core, front to back][tail H][tail V]. */
    const long alen = headH + headV + nops + tailH + tailV;
    char *rrow = (char *)malloc((size_t)alen + 1), *arow = (char *)malloc((size_t)alen + 1);
    if (!rrow || !arow) { free(rrow); free(arow); free(h); free(v); free(colM); free(colH); free(trace); free(ops); return -1; }
    {
        static const char letters[5] = { 'A', 'C', 'G', 'T', 'N' };
        long p = 0, ri = 0, ai = 0;
        for (long k = 0; k < headH; ++k) { rrow[p] = letters[h[ri++]]; arow[p] = '-'; ++p; }
        for (long k = 0; k < headV; ++k) { rrow[p] = '-'; arow[p] = letters[v[ai++]]; ++p; }
        if (path_meta) { path_meta[0] = headH; path_meta[1] = headV; path_meta[2] = tailH; path_meta[3] = tailV; path_meta[4] = best; }
        if (path_n) *path_n = nops;
        for (long k = nops - 1; k >= 0; --k) {
            if (path_ops) { path_ops[nops - 1 - k] = ops[k]; path_flags[nops - 1 - k] = (char)(ops[k] == 'D' && h[ri] == v[ai]); }
            if (ops[k] == 'D')      { rrow[p] = letters[h[ri++]]; arow[p] = letters[v[ai++]]; }
            else if (ops[k] == 'V') { rrow[p] = '-';              arow[p] = letters[v[ai++]]; }
            else                    { rrow[p] = letters[h[ri++]]; arow[p] = '-'; }
            ++p;
        }
        for (long k = 0; k < tailH; ++k) { rrow[p] = letters[h[ri++]]; arow[p] = '-'; ++p; }
        for (long k = 0; k < tailV; ++k) { rrow[p] = '-'; arow[p] = letters[v[ai++]]; ++p; }
        rrow[p] = arow[p] = '\0';
    }

    /* ---- ScoredAlignment (src/alignment.cpp:5-110), scan for scan */
    out->read_start = -1; out->adapter_start = -1; out->raw_score = best;
    out->read_end = 0; out->adapter_end = 0;
    out->aligned_matches = out->aligned_len = out->full_matches = out->full_len = 0;
    {
        long startPos = -1, endPos = -1;
        int rs = 0, as = 0;
        for (long i = 0; i < alen; ++i) {                       /* :26-37 */
            if (rrow[i] != '-') rs = 1;
            if (arow[i] != '-') as = 1;
            if (rs && as) { startPos = i; break; }
        }
        rs = as = 0;
        for (long i = alen - 1; i >= 0; --i) {                  /* :40-51 */
            if (rrow[i] != '-') rs = 1;
            if (arow[i] != '-') as = 1;
            if (rs && as) { endPos = i; break; }
        }
        if (startPos != -1 && endPos != -1) {                   /* :53-54 */
            long aStart = -1, aEnd = -1;
            for (long i = 0; i < alen; ++i) if (arow[i] != '-') { aStart = i; break; }        /* :57-63 */
            for (long i = alen - 1; i >= 0; --i) if (arow[i] != '-') { aEnd = i; break; }     /* :64-70 */
            int cnt = 0;
            for (long i = startPos; i < endPos + 1; ++i) if (arow[i] == rrow[i]) ++cnt;        /* :74-80 */
            out->aligned_matches = cnt; out->aligned_len = (int)(endPos - startPos + 1);
            cnt = 0;
            for (long i = aStart; i < aEnd + 1; ++i) if (arow[i] == rrow[i]) ++cnt;            /* :82-88 */
            out->full_matches = cnt; out->full_len = (int)(aEnd - aStart + 1);
            int rb = 0, ab = 0;
            for (long i = 0; i < alen; ++i) {                                                  /* :90-108 */
                if (i == startPos) { out->read_start = rb; out->adapter_start = ab; }
                if (i == endPos)   { out->read_end = rb;   out->adapter_end = ab; }
                if (rrow[i] != '-') ++rb;
                if (arow[i] != '-') ++ab;
            }
        }
    }
    free(rrow); free(arow); free(h); free(v); free(colM); free(colH); free(trace); free(ops);
    return 0;
}

int pc_oracle_align_n(const char *read_seq, long m, const char *adapter_seq, long n,
                      int match, int mismatch, int gap_open, int gap_extend, pc_oracle_result *out)
{
    return align_impl(read_seq, m, adapter_seq, n, match, mismatch, gap_open, gap_extend, out, NULL, NULL, NULL, NULL);
}

long pc_oracle_path(const char *read_seq, long m, const char *adapter_seq, long n,
                    int match, int mismatch, int gap_open, int gap_extend,
                    char *ops, char *flags, long *meta)
{
    pc_oracle_result r; long k = 0;
    if (align_impl(read_seq, m, adapter_seq, n, match, mismatch, gap_open, gap_extend, &r, ops, flags, meta, &k) != 0) return -1;
    return k;
}

int pc_oracle_align(const char *read_seq, const char *adapter_seq,
                    int match, int mismatch, int gap_open, int gap_extend, pc_oracle_result *out)
{
    return pc_oracle_align_n(read_seq, (long)strlen(read_seq), adapter_seq, (long)strlen(adapter_seq),
                             match, mismatch, gap_open, gap_extend, out);
}

void pc_oracle_format(const pc_oracle_result *r, char *buf, unsigned long buflen)
{
    /* getString(): std::to_string on ints and doubles = "%d" and "%f" (src/alignment.cpp:112-120);
     * identities as alignment.cpp:81,89: 100.0 * count / length in double. */
    double aligned = 0.0, full = 0.0;
    if (r->read_start != -1) {
        aligned = 100.0 * r->aligned_matches / r->aligned_len;
        full = 100.0 * r->full_matches / r->full_len;
    }
    snprintf(buf, buflen, "%d,%d,%d,%d,%d,%f,%f", r->read_start, r->read_end, r->adapter_start,
             r->adapter_end, r->raw_score, aligned, full);
}

char *pc_oracle_adapter_alignment(const char *read_seq, const char *adapter_seq,
                                  int match, int mismatch, int gap_open, int gap_extend)
{
    pc_oracle_result r;
    char *buf = (char *)malloc(160);
    if (!buf) return NULL;
    if (pc_oracle_align(read_seq, adapter_seq, match, mismatch, gap_open, gap_extend, &r) != 0) { free(buf); return NULL; }
    pc_oracle_format(&r, buf, 160);
    return buf;
}

void pc_oracle_free(char *p) { free(p); }

int pc_oracle_align_batch(const char *read_blob, const long *read_off,
                          const char *adapter_blob, const long *adapter_off,
                          const int *task_read, const long *task_start, const long *task_len,
                          const int *task_adapter, long n_tasks,
                          int match, int mismatch, int gap_open, int gap_extend,
                          pc_oracle_result *out)
{
    for (long t = 0; t < n_tasks; ++t) {
        const long ro = read_off[task_read[t]];
        const long ao = adapter_off[task_adapter[t]];
        const long al = adapter_off[task_adapter[t] + 1] - ao;
        if (pc_oracle_align_n(read_blob + ro + task_start[t], task_len[t], adapter_blob + ao, al,
                              match, mismatch, gap_open, gap_extend, &out[t]) != 0)
            return -1;
    }
    return 0;
}
