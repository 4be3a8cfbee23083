/*
 * porechop_oracle.h -- CPU oracle for the adapter-alignment hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (custom_porechop_abi_b200/,
 * include/) may include, link or call this.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs use it, and only as the checker or
 * the timed CPU baseline.
 *
 * It is a plain-C restatement of what the reference computes for ONE (read, adapter)
 * pair: SeqAn 2.4.0 globalAlignment with AlignConfig<true,true,true,true>,
 * Score<int,Simple>, SingleTrace/GapsLeft traceback
 *   (reference: porechop_abi/src/adapter_align.cpp:11-31)
 * followed by the ScoredAlignment post-processing
 *   (reference: porechop_abi/src/alignment.cpp:5-120).
 *
 * Parity status: PINNED.  tests/test_oracle_vs_reference.py checks it, on random and
 * adversarial inputs, against oracle/_ref/cpp_functions_ref.so (the reference's own two
 * .cpp files compiled where they lie, see oracle/Makefile) and against the committed
 * golden vectors under tests/golden/ that were generated with that library.
 */
#ifndef PORECHOP_ORACLE_H
#define PORECHOP_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pc_oracle_result {
    int read_start;       /* m_readStartPos    (-1: alignment failed, nothing else valid) */
    int read_end;         /* m_readEndPos      (inclusive, as the C++ prints it)          */
    int adapter_start;    /* m_adapterStartPos                                             */
    int adapter_end;      /* m_adapterEndPos                                               */
    int raw_score;        /* m_rawScore                                                    */
    int aligned_matches;  /* alignedMatchCount                                             */
    int aligned_len;      /* alignedRegionLength                                           */
    int full_matches;     /* fullAdapterMatchCount                                         */
    int full_len;         /* fullAdapterLength                                             */
} pc_oracle_result;

/* Runs the alignment; returns 0 on success, -1 on allocation failure. */
int pc_oracle_align(const char *read_seq, const char *adapter_seq,
                    int match, int mismatch, int gap_open, int gap_extend,
                    pc_oracle_result *out);

/* Same, on explicit lengths (no NUL needed). */
int pc_oracle_align_n(const char *read_seq, long read_len, const char *adapter_seq, long adapter_len,
                      int match, int mismatch, int gap_open, int gap_extend,
                      pc_oracle_result *out);

/* Test helper: the alignment path itself. ops/flags need read_len + adapter_len + 1 bytes; meta = 5 longs
 * {head_h, head_v, tail_h, tail_v, score}. Returns the number of core columns, -1 on failure. */
long pc_oracle_path(const char *read_seq, long read_len, const char *adapter_seq, long adapter_len,
                    int match, int mismatch, int gap_open, int gap_extend,
                    char *ops, char *flags, long *meta);

/* Formats a result exactly as ScoredAlignment::getString() (alignment.cpp:112-120). buf >= 160 bytes. */
void pc_oracle_format(const pc_oracle_result *r, char *buf, unsigned long buflen);

/* Drop-in twin of the reference entry point: malloc'd string, free with pc_oracle_free. */
char *pc_oracle_adapter_alignment(const char *read_seq, const char *adapter_seq,
                                  int match, int mismatch, int gap_open, int gap_extend);
void pc_oracle_free(char *p);

/* Batch helper for the timed CPU baseline: aligns task t = (reads[read_idx[t]] window, adapters[adapter_idx[t]]).
 * Sequences are given as one blob + offsets (offsets has n+1 entries). */
int pc_oracle_align_batch(const char *read_blob, const long *read_off,
                          const char *adapter_blob, const long *adapter_off,
                          const int *task_read, const long *task_start, const long *task_len,
                          const int *task_adapter, long n_tasks,
                          int match, int mismatch, int gap_open, int gap_extend,
                          pc_oracle_result *out);

#ifdef __cplusplus
}
#endif
#endif
