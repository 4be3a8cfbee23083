/*
 * pcb_kernels.h -- host-visible launch interface of the sm_100a kernels (internal to the library).
 */
#ifndef PCB_KERNELS_H
#define PCB_KERNELS_H

#include <stdint.h>
#include <cuda_runtime.h>
#include "../../include/porechop_b200.h"

#define PCB_RT          32        /* adapter rows held in registers per stripe                    */
#define PCB_BLOCK       128       /* threads per CTA (4 warps = 4 work units in flight)           */
#ifndef PCB_MIN_BLOCKS
#define PCB_MIN_BLOCKS  4         /* CTAs per SM the register allocation is held to               */
#endif
#define PCB_CODE_SCALE  8         /* arena bytes = Dna5 code * 8 (a byte shift for the fast kernels) */
#define PCB_NEG_INF     (-8192)   /* "minus infinity" in a signed 16-bit lane                     */
#define PCB_MAX_TRACE_M 8000      /* widest segment the trace kernel takes (order key is 14 bit)  */

/* fast kernels (pcb_fast.cu): affine schemes, adapters of <= 72 bases.  An adapter's kernel class picks the number of
 * adapter rows the kernel holds in registers (NR) and the checkpoint distance / flag-block width of the two-pass trace (C):
 *   class 0: exactly 24 bases (almost every Porechop adapter and barcode)   NR 24, exact
 *   class 1: 1..23 bases      NR 24      class 2: 25..32 bases   NR 32
 *   class 3: 33..40 bases     NR 40      class 4: 41..52 bases   NR 52
 *   class 5: 53..72 bases     NR 72   (the native "full sequence" barcode sets, adapters.py:466-476: 68 and 63 bases)
 *   class 6: 73..112 bases    NR 56 x two stripes: one task per thread and pass instead of two, the second 56 rows in the
 *            high halves one column behind the first (the rapid "full sequence" barcode sets, adapters.py:479-498: 102, 111)
 *   class 7: 113..127 bases   NR 72 x two stripes
 *   class PCF_NCLS: everything else -> general kernels (pcb_kernels.cu) */
#define PCF_NCLS          8
#define PCF_CLS_TWO       6       /* first two-stripe class */
#define PCF_MAX_N         72      /* rows of the one-stripe classes */
#define PCF_MAX_N2        127     /* rows of the two-stripe class (the position word has 7 row bits) */
/* a traceback's position word: row | column << 7 | state << 17 | gap openings << 20; its end cell: row | column << 7 |
 * (raw score + 2048) << 17 */
#define PCF_ROW_BITS      7
#define PCF_COL_BITS      10
#define PCF_MAX_TRACE_M   1000    /* widest segment of the two-pass trace kernel (no bias re-basing) */
#define PCF_MAX_SCAN_M    12000   /* widest scan chunk (order key is 14 bit)                      */
#ifndef PCF_C_SMALL
#define PCF_C_SMALL       8       /* C for NR <= 32 */
#endif
#ifndef PCF_C_LARGE
#define PCF_C_LARGE       8       /* C for NR > 32  */
#endif
static inline int pcf_class_nr(int cls) { return cls <= 1 ? 24 : (cls == 2 ? 32 : (cls == 3 ? 40 : (cls == 4 ? 52 : (cls == 6 ? 56 : 72)))); }
static inline int pcf_class_two(int cls) { return cls >= PCF_CLS_TWO && cls < PCF_NCLS; }
static inline int pcf_class_rows(int cls) { return pcf_class_two(cls) ? 2 * pcf_class_nr(cls) : pcf_class_nr(cls); }   /* adapter rows a thread computes per column */
static inline int pcf_class_c(int cls) { return pcf_class_nr(cls) <= 32 ? PCF_C_SMALL : PCF_C_LARGE; }
static inline int pcf_class_of_len(long len)
{
    if (len < 1 || len > PCF_MAX_N2) return PCF_NCLS;
    if (len > PCF_MAX_N) return len <= 112 ? 6 : 7;
    return len == 24 ? 0 : (len < 24 ? 1 : (len <= 32 ? 2 : (len <= 40 ? 3 : (len <= 52 ? 4 : 5))));
}

/* One alignment task as the kernels see it. 32 bytes. */
struct PcbTask {
    uint64_t off;        /* offset of the segment's first base in the code arena                  */
    int32_t  len;        /* columns in the segment (0 = empty slot)                               */
    int32_t  col0;       /* read column the segment starts at (segment = read[col0 .. col0+len))  */
    int32_t  m_true;     /* full read length (for tail / coordinates)                             */
    int32_t  forced_ord; /* >= 0: trace back from this order key only (second phase of long reads)*/
    int32_t  out_idx;    /* slot in the result array                                              */
    int32_t  own0;       /* scan only: first owned column (1-based, segment-local) - 1 = warm-up  */
};

/* A work unit = one warp = up to 64 tasks sharing one adapter. */
struct PcbUnit {
    int32_t adapter;
    int32_t ntask;
    int64_t task0;
};

struct PcbScores {
    int32_t match, mismatch, gap_open, gap_extend;
    int32_t linear;      /* gap_open == gap_extend: NeedlemanWunsch tie-breaking in the traceback */
};

/* Where the tasks of a launch come from. */
struct PcbSource {
    /* explicit: */
    const PcbUnit *units;        /* may be NULL -> cross-product mode                            */
    const PcbTask *tasks;
    /* cross product windows x ids: */
    int64_t  n_windows;
    int32_t  stride;
    const int32_t *win_len;
    const int32_t *ids;          /* adapters of this launch                                      */
    int32_t  n_ids;
    const int32_t *id_pos;       /* position of ids[k] in the caller's id list (NULL: k)         */
    int32_t  n_ids_total;        /* length of the caller's id list: out[w * n_ids_total + pos]   */
    int64_t  n_units;
};

struct PcbAdapters {
    const uint8_t *codes;        /* all adapters, Dna5 codes                                     */
    const int64_t *off;          /* n_adapters + 1                                               */
    int32_t  max_len;
    int32_t  n;                  /* adapters in the set                                          */
};

/* Work item of pass 2 of the fast trace (pcb_fast.cu): one traceback in progress.  32 bytes = one sector. */
struct PcfItem {
    uint32_t ref;        /* (unit - unit0) << 6 | task slot in the unit                                   */
    uint32_t pos;        /* where the walk stands: row | column << 7 | state << 17 | gap openings << 20   */
                         /* (state 4: still to be decided from V == M / H == M of the end cell)            */
    uint32_t endc;       /* end cell: row | column << 7 | (raw score + 2048) << 17                         */
    uint32_t a0, a1, a2; /* the walk's accumulators, see trace_finish                                      */
    uint32_t pad0, pad1;
};

/* scan (score-only) output per task slot */
struct PcbScanBest {
    int32_t key;      /* (score << 16) | ((0x3FFF - ord_local) << 2) | eqV << 1 | eqH ; INT_MIN = none */
};

struct PcbLaunch {
    PcbSource   src;
    PcbAdapters ad;
    PcbScores   sc;
    const uint8_t *arena;
    uint64_t arena_len;
    pcb_result *out;             /* trace kernel                                                  */
    int32_t    *scan_out;        /* scan kernel: one key per task (indexed by out_idx)            */
    uint8_t    *aux;             /* general trace kernel, may be NULL: 1 = the last column of the aligned region is a match */
    uint32_t   *scratch;         /* per-thread-slot scratch, see kernels                          */
    int32_t     mmax;            /* scratch sized for segments up to this many columns            */
    int32_t     smax;            /* ... and adapters up to this many stripes                      */
    int32_t    *err_flag;        /* device int: bit0 = a path ran into the window's left edge     */
    int32_t     grid;
    /* two-kernel trace of the fast classes (pcb_fast.cu): this launch covers units [unit0, unit1) of src; pass 1 leaves
     * every task's checkpoints in scratch ([unit - unit0][lane][pcf_ck_words]) and its end cell in entries */
    int64_t     unit0, unit1;
    uint4      *entries;         /* [unit - unit0][lane]: {end cell, score} of the lane's two tasks */
    PcfItem    *sorted, *surv;   /* pass 2: items in processing order; items that go on to the next generation */
    int32_t    *meta;            /* pass 2: bucket counts, bases, chunk table (pcf_meta_words)      */
    int32_t     n_major;         /* pass 2: majors = adapters of the launch (cross product: its id list; task lists: the adapter set) */
    int32_t     gen;             /* pass 2: generation 0 = first block of every traceback, 1, 2 = later blocks */
    int32_t     finish;          /* pass 2: the last generation (rounds keep their items until every path is complete) */
    int64_t     n_out;           /* records in `out` (index checks of the -DPCF_CHECKED build)       */
};

/* words of scratch per thread slot */
static inline uint64_t pcb_scratch_words_trace(int mmax, int smax)
{
    return (uint64_t)smax * mmax * (PCB_RT / 4)      /* trace flags      */
         + 2ull * (mmax + 1)                          /* stripe boundary  */
         + 4ull * smax * PCB_RT;                      /* final-column dumps (2 areas x M,H) */
}
static inline uint64_t pcb_scratch_words_scan(int mmax, int smax)
{
    (void)smax;
    return 2ull * (mmax + 1) + 4ull * smax * PCB_RT;
}

/* A long read of the explicit-task path: scanned in column chunks, then traced on one window. */
struct PcbLong {
    int64_t off;         /* first base of the read in the task arena                               */
    int32_t len;         /* read length                                                            */
    int32_t warm, span;  /* margins for this adapter length and scoring scheme                     */
    int32_t chunk;       /* owned columns per scan chunk                                           */
    int32_t chunk0;      /* index of its first chunk key in scan_out                               */
    int32_t nchunks;
    int32_t out_idx;     /* result slot                                                            */
    int32_t win_slot;    /* index of its trace-window task in the launch's task array              */
};
/* Picks each long read's best end cell from its chunk keys (score, then SeqAn's visiting order) and writes the
 * trace-window task for it. */
cudaError_t pcb_launch_merge(const PcbLong *longs, int64_t n_long, const int32_t *keys, PcbTask *wins, int32_t *err_flag,
                             cudaStream_t st);

/* ---- device-side planner of the explicit-task path ---------------------------------------------------------------
 * The caller's task list (segment offset, length, adapter id) is turned into kernel task arrays ON THE DEVICE: the host
 * only sees a histogram.  A task is `direct` (short segment: one trace task) or `long` (column chunks for the scan kernel
 * + one trace-window task, filled in after the scan by the merge kernel).  Items are bucketed by (kind, adapter,
 * length bucket) so that a warp's 64 tasks share the adapter and have similar lengths. */
#define PCB_LB        64                /* length buckets per adapter                                 */
#define PCB_KINDS     3                 /* 0 direct, 1 scan chunk, 2 trace window                     */
struct PcbAdInfo {
    int32_t n;            /* adapter length                                                          */
    int32_t cls;          /* kernel class of direct tasks and scan chunks                            */
    int32_t warm, span;   /* margins; valid when chunked != 0                  */
    int32_t direct_max;   /* segments up to this length are direct tasks                             */
    int32_t chunked;      /* 0: every segment is a direct task (margins not available)               */
};
struct PcbPlanArgs {
    const int64_t *seg_off; const int32_t *seg_len; const int32_t *seg_ad; int64_t n_tasks;
    const PcbAdInfo *ad; int32_t n_adapters; int32_t chunk;       /* chunk: target owned columns per scan chunk */
    int32_t *count;             /* [PCB_KINDS][n_adapters][PCB_LB] items per bucket (count pass)          */
    int32_t *mmax;              /* [PCB_KINDS][PCF_NCLS + 1] widest item per (kind, class)  (count pass)  */
    const int64_t *base;        /* [PCB_KINDS][n_adapters][PCB_LB] first slot of a bucket in `tasks`      */
    int32_t *cursor;            /* same shape, zeroed: next free slot of a bucket                          */
    PcbTask *tasks; PcbLong *longs;
    int64_t scan_task0, win_task0;   /* where the scan chunks / window tasks start in `tasks`              */
    int64_t end_lo, end_hi;          /* only tasks whose segment ends in (end_lo, end_hi] (one piece of the blob) */
};
cudaError_t pcb_launch_plan(const PcbPlanArgs &a, bool fill, cudaStream_t st);

/* reductions of the cross-product records (pcb_reduce.cu) */
cudaError_t pcb_launch_best(const pcb_result *rec, int64_t n_windows, int32_t n_ids, unsigned long long *keys, pcb_best *out, cudaStream_t st);
cudaError_t pcb_launch_hits(const pcb_result *rec, int64_t n_windows, int32_t n_ids, const pcb_hit_filter &f, int32_t *cnt, int32_t *bsum,
                            long long *total, pcb_hit *hits, int64_t cap, const int32_t *ident_col, int32_t n_want, uint16_t *ident,
                            int32_t *err_flag, cudaStream_t st);

cudaError_t pcb_launch_trims(const pcb_result *rec, int64_t n_windows, int32_t n_ids, const pcb_hit_filter &f, int32_t *trim, cudaStream_t st);
cudaError_t pcb_launch_top2(const uint16_t *ident, int64_t n_windows, int32_t n_want, pcb_top2 *out, cudaStream_t st);

cudaError_t pcb_launch_encode(const char *ascii, uint8_t *codes, uint64_t n, cudaStream_t st);
cudaError_t pcb_launch_gather_windows(const uint8_t *arena_t, const int64_t *roff, const int32_t *rlen, int64_t n, int end_size, int at_end,
                                      int stride, uint8_t *arena, int32_t *win_len, cudaStream_t st);
cudaError_t pcb_launch_trace(const PcbLaunch &L, cudaStream_t st);
cudaError_t pcb_launch_scan(const PcbLaunch &L, cudaStream_t st);
int pcb_occupancy_grid(int which /*0 trace, 1 scan*/);

/* fast kernels, by class (see above).  pcf_launch_scan: score-only keys.  pcf_launch_pass1 / pass2: the two parts of the
 * trace of the units [L.unit0, L.unit1), whose checkpoints must fit L.scratch (pcf_unit_words_trace per unit) */
cudaError_t pcf_launch_scan(const PcbLaunch &L, int cls, cudaStream_t st);
cudaError_t pcf_launch_pass1(const PcbLaunch &L, int cls, cudaStream_t st);
cudaError_t pcf_launch_pass2(const PcbLaunch &L, int cls, int grid_rounds, cudaStream_t st);
int pcf_occupancy_grid(int kind /* 0 scan, 1 pass 1, 2 rounds */, int cls);
int pcf_generations(int cls);       /* generations of pass 2 (sorted lists of tracebacks) */
int pcf_trace_launches(int cls);    /* kernel launches of one pass 1 + pass 2 */
/* Trace of the fast classes over a batch of units (pcf_launch_trace): pass 1 (score only) leaves the DP column of every
 * task at every C-th column in scratch (per unit and checkpoint NR/8 warp-wide 1 KB lines: 8 rows per thread, 16 bits per
 * cell and task = M in 13 bits, M - H clamped in 3)
 * and each task's end cell in `entries`; pass 2 recomputes, with trace flags, only the C-column blocks a path crosses:
 * the tracebacks are sorted by what their next block costs (columns x rows), 64 of a kind make a round. */
#define PCF_NB 32                 /* sort buckets per major (first blocks: by width / rows; later blocks: by rows) */
static inline uint64_t pcf_ck_words(int mmax, int C, int NR) { return (uint64_t)((mmax + C - 1) / C) * (uint64_t)((NR + 7) / 8) * 8ull; }
/* words per unit: checkpoints, entries (4 per thread), sorted items and surviving items (8 words per task each) */
/* two-stripe class: 2 tasks x checkpoints (one more iteration than columns) x (row chunks + 1 chunk of boundary registers) */
static inline uint64_t pcf_ck_words2(int mmax, int C, int NR) { return 2ull * (uint64_t)((mmax + 1 + C - 1) / C) * (uint64_t)((NR + 7) / 8 + 1) * 8ull; }
static inline uint64_t pcf_unit_words_trace(int mmax, int cls)
{
    const uint64_t ck = pcf_class_two(cls) ? pcf_ck_words2(mmax, pcf_class_c(cls), pcf_class_nr(cls)) : pcf_ck_words(mmax, pcf_class_c(cls), pcf_class_nr(cls));
    return 32ull * (ck + 4) + 2ull * 64 * 8;
}
/* meta (int32): hist[2][A * NB] | base[A * NB + 1] | cursor[A * NB] | chunk_base[A + 1] | counters[8] */
static inline uint64_t pcf_meta_words(int n_major) { return 4ull * n_major * PCF_NB + n_major + 16; }

#endif
