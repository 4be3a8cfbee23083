/*
 * porechop_b200.h -- C ABI of the B200-native adapter-alignment library (cpp_functions.so).
 *
 * This is the drop-in boundary.  The first two entry points are EXACTLY what the reference's
 * ctypes wrapper binds (reference: porechop_abi/cpp_function_wrappers.py:27-39, declared in
 * porechop_abi/include/adapter_align.h:12-16, implemented in porechop_abi/src/adapter_align.cpp:11-36).
 * The pcb_* entry points are the batched form of the same call: the reference makes one
 * adapterAlignment call per (sequence window, adapter) pair from
 *   nanopore_read.py:158-173 (align_adapter_set), :175-217 (find_start_trim / find_end_trim),
 *   :219-252 (find_middle_adapters);
 * the batched entry points take all those pairs at once and return, per pair, the integers the
 * result string is printed from (porechop_abi/src/alignment.cpp:112-120).
 *
 * Plain pointers and sizes only; no torch / CUDA types appear in any signature.  All compute runs
 * in hand-written sm_100a CUDA kernels; there is no CPU fallback: every compute entry point fails
 * (non-zero return, pcb_last_error() says why) when no CUDA device is usable.
 */
#ifndef PORECHOP_B200_H
#define PORECHOP_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- 1. the reference's own two symbols (adapter_align.h:12-16) ------------------------------ */

/* Returns a malloc'd "readStart,readEnd,adapterStart,adapterEnd,rawScore,alignedId,fullId" string
 * (alignment.cpp:112-120), to be released with freeCString.  The reference's function cannot fail on
 * non-empty input and has no error channel, so a library failure (no device, CUDA error, scoring scheme
 * outside the 16-bit range) ends the process with the reason on stderr -- an unmodified caller would
 * otherwise read the failure record as "no adapter" and write wrong output.  After
 * pcb_set_dropin_fatal(0) the call instead returns the reference's failure encoding (readStart == -1,
 * alignment.cpp:7-21) and pcb_last_error(NULL) holds the reason (per thread). */
char *adapterAlignment(char *readSeq, char *adapterSeq,
                       int matchScore, int mismatchScore, int gapOpenScore, int gapExtensionScore);
void freeCString(char *p);
int pcb_set_dropin_fatal(int on);

/* ---- 2. batched form ------------------------------------------------------------------------- */

typedef struct pcb_context pcb_context;

/* The integers ScoredAlignment holds (alignment.h:17-25).  alignedRegionPercentIdentity is
 * 100.0 * matches / aligned_len and fullAdapterPercentIdentity is 100.0 * matches / full_len,
 * evaluated in double exactly as alignment.cpp:81,89 do (the two match counts of the reference are
 * always equal).  read_start == -1 marks the reference's "alignment failed" case
 * (empty sequence). read_end is inclusive, as the C++ prints it. */
typedef struct pcb_result {
    int32_t read_start;
    int32_t read_end;
    int32_t adapter_start;
    int32_t adapter_end;
    int32_t raw_score;
    int32_t matches;
    int32_t aligned_len;
    int32_t full_len;
} pcb_result;

/* Error codes returned by the pcb_* calls (0 = ok). */
#define PCB_OK            0
#define PCB_ERR_CUDA      1   /* a CUDA runtime call failed or no device is present       */
#define PCB_ERR_ARG       2   /* bad argument                                              */
#define PCB_ERR_RANGE     3   /* scoring scheme / lengths outside what the kernels support */
#define PCB_ERR_NOMEM     4

int  pcb_create(int device, pcb_context **out);
void pcb_destroy(pcb_context *ctx);
/* Last error text of this context (ctx == NULL: of the process-wide context adapterAlignment uses). */
const char *pcb_last_error(pcb_context *ctx);
int  pcb_device_count(void);

/* Scoring scheme, argument order as adapterAlignment's. */
int pcb_set_scoring(pcb_context *ctx, int match, int mismatch, int gap_open, int gap_extend);

/* Adapter set: ASCII blob + n_adapters+1 offsets.  Converted with SeqAn's Dna5 table
 * (seqan/basic/alphabet_residue_tabs.h:113-140) on the device. */
int pcb_set_adapters(pcb_context *ctx, const char *blob, const int64_t *offsets, int32_t n_adapters);

/* Cross product "every window x every listed adapter" (phases A and B of the reference:
 * porechop_abi.py:200-245, :359-438).  windows: n_windows records of `stride` bytes each, of which
 * the first win_len[w] bytes are the window (ASCII).  out[w * n_ids + k] = window w vs adapter_ids[k].
 * HOST buffers; copies are part of the call. */
int pcb_align_windows(pcb_context *ctx, const char *windows, int64_t n_windows, int32_t stride,
                      const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids,
                      pcb_result *out);

/* ---- 2b. the cross product, reduced on the device to what the reference's callers consume ---------------------
 * pcb_align_windows returns 32 bytes per (window, adapter) pair; the two calls below run the same kernels and hand back
 * a few percent of that. */

/* Phase A (porechop_abi.py:200-245): NanoporeRead.align_adapter_set keeps, per adapter sequence, the MAXIMUM
 * fullAdapterPercentIdentity over all check reads (nanopore_read.py:158-173).  best[k] = (matches, full_len) of a pair
 * with the largest 100 * matches / full_len for adapter_ids[k]; (0, 0) when no window aligned (identity 0.0).  Ratios
 * are compared exactly (as integers); the printed "%f" of two different ratios of lengths < 4096 always differs. */
typedef struct pcb_best { int32_t matches, full_len; } pcb_best;
int pcb_align_windows_best(pcb_context *ctx, const char *windows, int64_t n_windows, int32_t stride,
                           const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids, pcb_best *best);

/* Phase B (porechop_abi.py:359-438): find_start_trim / find_end_trim keep a pair only if (nanopore_read.py:187-188,
 * :209-210)  partial_score > end_threshold  and  read_end != end_size (start windows) / read_start != 0 (end windows)
 * and  read_end - read_start >= min_trim_size,  where partial_score is the aligned-region identity as parsed from the
 * printed "%f" and read_end is exclusive.  threshold_e6 = floor(end_threshold * 10^6): the printed identity exceeds the
 * threshold exactly when 100 * matches / aligned_len >= (threshold_e6 + 1/2) * 10^-6. */
typedef struct pcb_hit_filter {
    int64_t threshold_e6;
    int32_t at_end;          /* 0: start windows, 1: end windows */
    int32_t end_size;
    int32_t min_trim_size;
    int32_t reserved;
} pcb_hit_filter;
/* A kept pair: window index, position of the adapter in adapter_ids, and the record's fields the callers read
 * (read_end inclusive, as in pcb_result). */
typedef struct pcb_hit {
    int32_t window, id_pos, read_start, read_end, raw_score, matches, aligned_len, full_len;
} pcb_hit;
/* hits: the kept pairs, ordered by window and then by id_pos (the order in which the reference appends them to
 * start_adapter_alignments / end_adapter_alignments); *n_hits = their number (more than hits_cap: the call fails with
 * PCB_ERR_RANGE after filling hits_cap of them).  want_ident (may be NULL): per adapter of the list 1 = also return its
 * full-adapter identity for EVERY window (barcode scores, nanopore_read.py:190-195, :212-217): ident[w * n_want + j]
 * for the j-th wanted adapter = matches << 9 | full_len (0: the alignment failed, identity 0.0); ident == NULL keeps
 * them on the device (for pcb_fetch_top2).  In deferred mode hits / *n_hits are valid after pcb_wait. */
int pcb_align_windows_hits(pcb_context *ctx, const char *windows, int64_t n_windows, int32_t stride,
                           const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids,
                           const pcb_hit_filter *filter, const uint8_t *want_ident,
                           pcb_hit *hits, int64_t hits_cap, int64_t *n_hits, uint16_t *ident);

/* End windows cut out of the announced read blob ON THE DEVICE: window r = the first (at_end = 0) or last (at_end = 1)
 * min(end_size, read_len[r]) bases of blob[read_off[r] .. +read_len[r]) (nanopore_read.py:164,170,181,203).  Needs
 * pcb_upload_reads; stages the selected slot like pcb_stage_windows, after which pcb_align_windows_best / _hits are called
 * with windows == NULL (n_windows = n_reads) and pcb_run_windows_resident works as usual.  The reads then cross PCIe once
 * for all three phases, and the host never gathers windows. */
int pcb_stage_windows_from_reads(pcb_context *ctx, const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                 int32_t end_size, int32_t at_end);

/* Barcode binning (nanopore_read.py:408-482): determine_barcode sorts a read's barcode scores in descending order
 * (stable: the first in adapter order among equals) and decides on the first two.  After a pcb_align_windows_hits call
 * with want_ident, top2[w] = positions (among the wanted adapters) and identity words of the best and second-best
 * barcode of window w (-1 / 0 when there are fewer); identities are compared exactly.  Same slot as the hits call;
 * in deferred mode valid after pcb_wait. */
typedef struct pcb_top2 { int16_t best, second; uint16_t best_ident, second_ident; } pcb_top2;
int pcb_fetch_top2(pcb_context *ctx, pcb_top2 *top2, int64_t n_windows);

/* The trim amount the kept alignments of window w ask for, before extra_trim_size is added (find_start_trim /
 * find_end_trim, nanopore_read.py:187-190, :209-212: the maximum of read_end at the start, of end_size - read_start at the
 * end), -1 when none was kept: what phase C needs of phase B, 4 bytes per read end.  Same slot and filter as the
 * pcb_align_windows_hits call before it; in deferred mode valid after pcb_wait_slot / pcb_wait. */
int pcb_fetch_trims(pcb_context *ctx, int32_t *trim, int64_t n_windows);

/* Explicit task list (phase C, porechop_abi.py:457-522, and anything irregular): task t aligns
 * seq_blob[seg_off[t] .. seg_off[t]+seg_len[t]) against adapter task_adapter[t].  Any length.
 * HOST buffers. */
int pcb_align_tasks(pcb_context *ctx, const char *seq_blob, int64_t blob_len,
                    const int64_t *seg_off, const int32_t *seg_len, const int32_t *task_adapter,
                    int64_t n_tasks, pcb_result *out);

/* Announces the read blob of a later pcb_align_tasks call (the reference holds all reads in memory before the
 * first alignment, porechop_abi.py:61-71, so the blob is known while phases A and B still run).  Returns at
 * once; the pcb_align_windows calls that follow send the blob to the device, in pieces, under their own
 * kernels.  The matching call is pcb_align_tasks(ctx, NULL, blob_len, ...): a NULL seq_blob means "the
 * announced blob".  The host buffer (pinned, for the copies to overlap) must stay untouched until that call
 * returns.  A pcb_align_tasks / pcb_stage_tasks call with its own blob cancels the announcement. */
int pcb_upload_reads(pcb_context *ctx, const char *seq_blob, int64_t blob_len);

/* Submit / collect.  The reference queues its alignments on a thread pool and collects them when the pool has
 * drained (porechop_abi.py:225-231, :395-407, :489-499).  pcb_set_deferred(ctx, 1) gives the host-buffer calls the
 * same shape: pcb_align_windows / pcb_align_tasks return once their copies and kernels are queued, so the copies of
 * one call run under the kernels of the one before, and pcb_wait(ctx) returns when every queued call is complete,
 * its results are in the callers' buffers, and reports the first device-side error of the batch.  Between two
 * pcb_wait calls: input and output buffers stay untouched and every call uses its own slot (pcb_select_slot).  Several
 * pcb_align_tasks(ctx, NULL, ...) calls on one announced blob may be queued (a batch that goes through in parts): their
 * plans alternate between two sets of device arrays, so the planner of one runs beside the kernels of the one before.
 * pcb_set_deferred waits for queued work before it switches. */
int pcb_set_deferred(pcb_context *ctx, int on);
int pcb_wait(pcb_context *ctx);
/* Collects ONE slot's queued pcb_align_windows_hits call while the calls queued in other slots keep running: the caller
 * can derive that part's trims and queue its phase C beside the phase B of the next part (the reference's pool works
 * through reads in order in the same way).  Device-side errors are reported by the pcb_wait that ends the batch. */
int pcb_wait_slot(pcb_context *ctx, int slot);

/* ---- 2c. the other call site of the same DP: check_compatibility --------------------------------------------
 * porechop_abi/ab_initio_src/compatibility.h:56-58 / compatibility.cpp:160-168 (compatibility.so, bound by
 * consensus.py:72-100): semi-global alignment of two candidate adapters as String<Dna>, longer sequence first, linear gaps,
 * match 2 / mismatch -1 / gap -1; flag 0 = not compatible, 1 = compatible (integer identity of the aligned region >=
 * 87.5 %, errors counted without its last column, as the reference does), 2 = compatible and the shorter sequence lies
 * inside the longer one (more than 3 bases from either end).  check_compatibility has the reference's signature;
 * pcb_check_compatibility takes all sequences and a list of index pairs at once (the reference calls it all-vs-all).
 * Both replace the context's adapter set; sequences must not be empty. */
int check_compatibility(char *raw_seq1, char *raw_seq2);
int pcb_check_compatibility(pcb_context *ctx, const char *blob, const int64_t *offsets, int32_t n_seqs,
                            const int32_t *pair_a, const int32_t *pair_b, int64_t n_pairs, int32_t *flags);

/* ---- 3. device-resident form (same kernels; used to time the kernels without PCIe) ---------- */

/* A context holds 8 independent staging slots (inputs + results); all stage/run/fetch calls, and the host-buffer
 * calls above, act on the selected one (default 0).  Lets start windows, end windows and full-length tasks stay
 * resident side by side. */
int pcb_select_slot(pcb_context *ctx, int slot);

/* Stage inputs in HBM (H2D + Dna5 encode). */
int pcb_stage_windows(pcb_context *ctx, const char *windows, int64_t n_windows, int32_t stride,
                      const int32_t *win_len);
int pcb_stage_ids(pcb_context *ctx, const int32_t *adapter_ids, int32_t n_ids);
/* Run the cross product on the staged windows; results stay in HBM. */
int pcb_run_windows_resident(pcb_context *ctx);
/* Copy the staged run's results to the host (n_windows * n_ids records). */
int pcb_fetch_results(pcb_context *ctx, pcb_result *out, int64_t n_results);

int pcb_stage_tasks(pcb_context *ctx, const char *seq_blob, int64_t blob_len,
                    const int64_t *seg_off, const int32_t *seg_len, const int32_t *task_adapter,
                    int64_t n_tasks);
int pcb_run_tasks_resident(pcb_context *ctx);

/* CUDA-event timer on the library's own stream. */
int pcb_timer_start(pcb_context *ctx);
int pcb_timer_stop(pcb_context *ctx, float *ms);
int pcb_synchronize(pcb_context *ctx);

/* Counters since context creation: kernel launches, DP cell updates issued, H2D / D2H bytes. */
typedef struct pcb_counters {
    int64_t kernel_launches;
    int64_t dp_cells;
    int64_t h2d_bytes;
    int64_t d2h_bytes;
    int64_t dp_kernel_launches;
    float   dp_kernel_ms;      /* summed CUDA-event time of the DP kernels (when profiling enabled) */
} pcb_counters;
int pcb_get_counters(pcb_context *ctx, pcb_counters *out);
int pcb_reset_counters(pcb_context *ctx);
/* When on, every DP kernel launch is bracketed by its own event pair (adds sync at fetch time). */
int pcb_set_kernel_timing(pcb_context *ctx, int on);

/* ---- windows x one adapter without any allocation -------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Needs no pcb_context.
 * Same kernels and the same records as pcb_align_windows with a one-adapter list (the two-pass trace of the fast kernel
 * classes), but every byte of working memory -- encoded windows, checkpoints, work items, sort tables, error flag --
 * lives in one file-scope device array of PCB_STATIC_SCRATCH_MB that every GPU loading the library holds until the
 * process exits.  The call itself calls no cudaMalloc / cudaFree or allocating library routine.
 *   d_windows, d_win_len, d_out: DEVICE memory of the caller on `device`: n_windows records of `stride` ASCII bytes,
 *   their lengths (0..stride), and n_windows result records.  adapter: HOST string of adapter_len bases.
 *   cuda_stream: a cudaStream_t of that device (NULL = the default stream); the call returns when its work on that
 *   stream is complete, d_out is valid then.
 * Limits (PCB_ERR_RANGE with a text in pcb_last_error(NULL) beyond them): n_windows <= PCB_STATIC_MAX_WINDOWS, stride <=
 * PCB_STATIC_MAX_STRIDE, adapters of 1..127 bases of ACGT, affine schemes with -7 <= gap_open < gap_extend < 0 < match.
 * All calls share the array and find the previous call's contents in it; the call clears what must start at zero.
 * Calls that overlap on one GPU would race on it: the calls of one process are serialised by a lock that is held until
 * the call's stream work is done (every process has its own copy of the array). */
#define PCB_STATIC_SCRATCH_MB   16
#define PCB_STATIC_MAX_WINDOWS  8192
#define PCB_STATIC_MAX_STRIDE   256
int pcb_align_windows_static(int device, const char *adapter, int32_t adapter_len, int match, int mismatch,
                             int gap_open, int gap_extend, const char *d_windows, const int32_t *d_win_len,
                             int64_t n_windows, int32_t stride, pcb_result *d_out, void *cuda_stream);

/* Formats one record exactly as ScoredAlignment::getString(); buf >= 160 bytes. */
void pcb_format_result(const pcb_result *r, char *buf, uint64_t buflen);

/* Host-only helper (no device needed): the ScoredAlignment statistics from an explicit alignment
 * path.  ops = core path front to back, 'D' / 'V' (adapter base over read gap) / 'H'.  Exposed so that
 * the CPU test-suite can check the same code the kernels run for their statistics. */
int pcb_stats_from_path(const char *ops, int64_t n_ops, int64_t head_h, int64_t head_v,
                        int64_t tail_h, int64_t tail_v, const uint8_t *match_flags,
                        int32_t raw_score, pcb_result *out);

#ifdef __cplusplus
}
#endif
#endif
