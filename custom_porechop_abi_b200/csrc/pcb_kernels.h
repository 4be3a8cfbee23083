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

/* fast kernels (pcb_fast.cu): affine schemes, adapters of <= PCB_RT bases */
#define PCF_C             16      /* checkpoint distance = width of a flag block (columns)        */
#define PCF_MAX_TRACE_M   1000    /* widest segment of the two-pass trace kernel (no bias re-basing) */
#define PCF_MAX_SCAN_M    12000   /* widest scan chunk (order key is 14 bit)                      */

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
    uint32_t   *scratch;         /* per-thread-slot scratch, see kernels                          */
    int32_t     mmax;            /* scratch sized for segments up to this many columns            */
    int32_t     smax;            /* ... and adapters up to this many stripes                      */
    int32_t    *err_flag;        /* device int: bit0 = a path ran into the window's left edge     */
    int32_t     grid;
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

cudaError_t pcb_launch_encode(const char *ascii, uint8_t *codes, uint64_t n, cudaStream_t st);
cudaError_t pcb_launch_trace(const PcbLaunch &L, cudaStream_t st);
cudaError_t pcb_launch_scan(const PcbLaunch &L, cudaStream_t st);
int pcb_occupancy_grid(int which /*0 trace, 1 scan*/);

/* fast kernels: cls 0 = adapters of exactly 24 bases, cls 1 = adapters of 1..32 bases */
#define PCF_EXACT_N 24
cudaError_t pcf_launch(const PcbLaunch &L, bool trace, int cls, cudaStream_t st);
int pcf_occupancy_grid(bool trace, int cls);
static inline uint64_t pcf_scratch_words_trace(int mmax, int cls)
{
    return (uint64_t)((mmax + PCF_C - 1) / PCF_C + 1) * 2ull * (cls == 0 ? PCF_EXACT_N : PCB_RT);     /* checkpoints */
}

#endif
