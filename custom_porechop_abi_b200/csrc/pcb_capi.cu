/*
 * pcb_capi.cu -- host side of cpp_functions.so: the C ABI declared in include/porechop_b200.h.
 *
 * Replaces the reference's native library (porechop_abi/src/adapter_align.cpp, src/alignment.cpp) behind
 * the same two symbols its ctypes wrapper binds (porechop_abi/cpp_function_wrappers.py:27-39), and adds the
 * batched submit/collect entry points.  No CPU alignment code lives here: every result comes out of the
 * kernels in pcb_kernels.cu.  The only arithmetic done on the host is the final
 * 100.0 * matches / length in double (alignment.cpp:81,89) when a result string is printed.
 */
#include "pcb_kernels.h"
#include "align_stats.h"

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

namespace {

template <typename T>
struct DevBuf {
    T *p = nullptr;
    size_t cap = 0;      // elements
    cudaError_t reserve(size_t n)
    {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = n + n / 8 + 64;
        cudaError_t e = cudaMalloc((void **)&p, want * sizeof(T));
        if (e != cudaSuccess) { want = n; e = cudaMalloc((void **)&p, want * sizeof(T)); }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

}  // namespace

struct pcb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t cstream = nullptr;       // result copies (D2H) that run beside the kernels of `stream`
    cudaStream_t ustream = nullptr;       // read-blob uploads (H2D + encode) that run beside them
    int64_t uploaded_len = -1;            // pcb_upload_reads announced a blob of this length
    cudaStream_t hstream = nullptr;       // hit lists of single slots (pcb_wait_slot), beside everything else
    enum { N_SIDE = 3 };
    cudaStream_t sstream[N_SIDE] = {};    // cross product: the parts of a call run as independent pipelines side by side (run_windows)
    cudaEvent_t ev_fork = nullptr, ev_join[N_SIDE] = {};
    int n_pipes = 4;                      // pipelines a cross product is dealt to (main stream + side streams; PCB_STREAMS)
    cudaStream_t pstream = nullptr;       // deferred mode: task staging + planning, beside the kernels of earlier calls
    cudaEvent_t tasks_done[2] = {};       // the task kernels that ran on plan set 0 / 1 no longer read its arrays
    bool deferred = false;                // pcb_set_deferred: host-buffer calls only enqueue; pcb_wait collects
    cudaStream_t in_stream() const { return deferred ? ustream : stream; }     // window / id staging copies
    cudaStream_t plan_stream() const { return deferred ? pstream : stream; }   // segment copies + planner kernels
    cudaEvent_t pipe_ev[8] = {};
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, kev0 = nullptr, kev1 = nullptr;
    std::string err;
    PcbScores sc{3, -6, -5, -2, 0};
    // adapters
    std::vector<int64_t> ad_off;          // host copy
    std::vector<uint8_t> ad_fast;         // kernel class of an adapter (pcb_kernels.h): 0..PCF_NCLS-1 fast kernels (ACGT only), PCF_NCLS = general kernels
    uint64_t stamp = 1;                   // bumped when the scoring scheme or the adapter set changes
    int32_t n_adapters = 0;
    DevBuf<uint8_t> d_ad_codes;
    DevBuf<int64_t> d_ad_off;
    // staged inputs
    DevBuf<char> d_ascii;
    DevBuf<char> d_ascii_t;           // ASCII staging of the read blob (its own buffer: uploads run beside window staging)
    DevBuf<uint8_t> d_arena_t;        // explicit-task arena (kept apart so both inputs can stay resident)
    uint64_t arena_t_len = 0;
    // staged window sets + their results: a context holds PCB_SLOTS independent sets so that several inputs
    // (start windows, end windows, full-length tasks) can be resident at once; pcb_select_slot picks one
    struct Slot {
        DevBuf<uint8_t> d_arena;          // window arena
        DevBuf<char> d_ascii;             // its ASCII staging (per slot: in deferred mode the next call's copy runs beside this one's kernels)
        cudaEvent_t staged = nullptr;     // deferred mode: the slot's staging copies have arrived
        cudaEvent_t done = nullptr;       // deferred mode: the slot's last reduced call has delivered its fixed-size results (pcb_wait_slot)
        uint64_t arena_len = 0;
        int64_t sum_winlen = 0;
        DevBuf<int32_t> d_winlen, d_ids;      // d_ids: [fast ids | other ids | fast positions | other positions]
        std::vector<int32_t> h_ids, h_split;
        int32_t n_cls[PCF_NCLS + 1] = {}; uint64_t split_stamp = 0;
        int64_t n_windows = 0; int32_t stride = 0, n_ids = 0, win_mmax = 0;
        DevBuf<pcb_result> d_out;
        int64_t n_results = 0;
        // reduced outputs (pcb_align_windows_best / _hits)
        DevBuf<unsigned long long> d_keys;
        DevBuf<pcb_best> d_best;
        DevBuf<int32_t> d_cnt, d_identcol;
        DevBuf<pcb_hit> d_hits;
        DevBuf<uint16_t> d_ident;
        DevBuf<long long> d_total;
        DevBuf<pcb_top2> d_top2;
        DevBuf<int32_t> d_trim;           // pcb_fetch_trims
        pcb_hit_filter hit_filter{};      // filter of the slot's last pcb_align_windows_hits call
        bool has_filter = false;
        DevBuf<int64_t> d_roff;           // pcb_stage_windows_from_reads: where the reads are in the blob
        DevBuf<int32_t> d_rlen;
        DevBuf<uint8_t> d_aux;            // check_compatibility: end-column match of every task
        bool want_aux = false;
        int32_t n_want = 0;
        std::vector<int32_t> h_identcol;
    };
    enum { N_SLOTS = 8 };
    Slot slots[N_SLOTS];
    int cur = 0;
    Slot &S() { return slots[cur]; }
    // explicit tasks
    int64_t n_tasks = 0, task_long_cols = 0, task_cells = 0, task_max_end = 0;   // task_max_end: where the last segment of the staged list ends in the blob
    DevBuf<int64_t> d_seg_off, d_plan_i64;
    DevBuf<int32_t> d_seg_len, d_seg_ad, d_plan_i32;
    DevBuf<PcbAdInfo> d_adinfo;
    struct TaskPlan {
        DevBuf<PcbTask> d_tasks;
        DevBuf<PcbUnit> d_units;
        DevBuf<int32_t> d_scan;
        DevBuf<PcbLong> d_longs;
        int64_t b0 = 0, b1 = 0;              // byte range of the piece
        int64_t end_lo = -1, end_hi = 0;     // the piece of the blob this plan covers: tasks whose segment ends in (end_lo, end_hi]
        cudaEvent_t ready = nullptr;         // recorded on the copy stream once that piece is in the arena
        bool has_ready = false;
        bool encoded = true;          // false: the ASCII of [b0, b1) has been sent but not yet turned into codes
        void release() { d_tasks.release(); d_units.release(); d_scan.release(); d_longs.release(); }
        struct Launch { bool trace; int cls, kind /*0 direct, 1 scan, 2 window*/; int64_t cells, task0, unit0, n_units; int mmax, nmax, smax, grid; };
        std::vector<Launch> launches;
        int64_t n_long = 0, n_chunks = 0, win_task0 = 0;
        uint64_t scratch_words = 0, stamp = 0;
        bool valid = false;
    };
    enum { MAX_PIECES = 16 };   // most pieces a blob travels in (a blob that comes with its tasks gets a plan per piece: blob_pieces)
    TaskPlan plans[MAX_PIECES];
    // Two task calls on one announced blob (a step that goes through in parts): the second call's planner fills `alt_plan`
    // while the first call's kernels still read plans[0], and the other way round the next time (plan_flip); the pieces
    // of the blob are described by plans[] either way.
    TaskPlan alt_plan;
    int plan_flip = 0;
    TaskPlan &one_plan() { return plan_flip ? alt_plan : plans[0]; }
    int n_pieces = 1, pieces_sent = 0;
    int plan_pieces = 1;                 // plans over the blob: one per piece (kernels of piece p under the copy of piece p+1), or one for
                                         // the whole blob when it was announced early and is on the device before its tasks are (pcb_upload_reads)
    const char *pending_blob = nullptr;  // host blob whose later pieces are sent once the plans exist (small planning copies
                                         // would otherwise queue behind the whole blob on the copy engine)
    // outputs / scratch
    DevBuf<uint32_t> d_scratch;
    DevBuf<int32_t> d_meta;               // fast trace, pass 2: bucket counts and chunk table (pcf_meta_words)
    DevBuf<uint32_t> d_scratch_s[N_SIDE]; // the same two for the side streams
    DevBuf<int32_t> d_meta_s[N_SIDE];
    DevBuf<int32_t> d_err;
    int grid_trace = 0, grid_scan = 0, grid_fp1[PCF_NCLS] = {}, grid_fp2[PCF_NCLS] = {}, grid_fscan[PCF_NCLS] = {};   // fast kernels: pass 1, rounds, scan
    // words of checkpoint scratch the batches of the fast trace may use together (PCB_SCRATCH_MB).  16 GB of the 180: a class's
    // units then go through in one batch per pipeline where 6 GB made three (131,072 windows x 108 sequences: 13.00 / 12.42 /
    // 12.08 / 12.08 ms with 3 / 6 / 16 / 32 GB); the buffers grow on demand, small inputs take what they need
    uint64_t scratch_budget = 16384ull << 18;
    pcb_counters cnt{};
    bool ktiming = false;
    // deferred pcb_align_windows_hits calls: their hit lists are copied once the counts are known (pcb_wait)
    struct PendingHits { const pcb_hit *d_hits; pcb_hit *hits; int64_t cap; int64_t *n_hits; int idx; };
    std::vector<PendingHits> pending_hits;
    long long *h_totals = nullptr;        // pinned, one per slot

    int fail(int code, const char *fmt, ...)
    {
        char buf[512];
        va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
        err = buf;
        return code;
    }
    int cuda_fail(cudaError_t e, const char *what)
    {
        return fail(PCB_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e));
    }
};

#define PCB_CUDA(ctx, call)                                                                  \
    do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return (ctx)->cuda_fail(e__, #call); } while (0)

// Error text of the process-wide entry points (pcb_create failures, adapterAlignment): one per thread, so that a pool of
// threads calling adapterAlignment (the reference's ThreadPool, porechop_abi.py:25) cannot clear or tear each other's.
static thread_local std::string g_global_err;
// adapterAlignment has no error channel in the reference (its library cannot fail on non-empty input).  An unmodified
// caller would read a failure record as "no adapter here" and write wrong output, so by default a library failure
// inside the drop-in symbol ends the process with the reason on stderr, as the reference's wrapper does for a missing
// library (cpp_function_wrappers.py:23-25).  A caller that checks pcb_last_error(NULL) itself switches that off.
static bool g_dropin_fatal = true;
extern "C" int pcb_set_dropin_fatal(int on) { g_dropin_fatal = on != 0; return PCB_OK; }

extern "C" int pcb_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int pcb_create(int device, pcb_context **out)
{
    if (!out) return PCB_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        g_global_err = std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        cudaGetLastError();
        return PCB_ERR_CUDA;
    }
    if (device < 0 || device >= n) { g_global_err = "device index out of range"; return PCB_ERR_ARG; }
    pcb_context *c = new pcb_context();
    c->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&c->cstream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&c->ustream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaStreamCreateWithPriority(&c->pstream, cudaStreamNonBlocking, -1)) != cudaSuccess ||
        (e = cudaStreamCreateWithFlags(&c->hstream, cudaStreamNonBlocking)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->tasks_done[0], cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreateWithFlags(&c->tasks_done[1], cudaEventDisableTiming)) != cudaSuccess ||
        (e = cudaEventCreate(&c->ev0)) != cudaSuccess || (e = cudaEventCreate(&c->ev1)) != cudaSuccess ||
        (e = cudaEventCreate(&c->kev0)) != cudaSuccess || (e = cudaEventCreate(&c->kev1)) != cudaSuccess) {
        g_global_err = std::string("context setup failed: ") + cudaGetErrorString(e);
        delete c;
        return PCB_ERR_CUDA;
    }
    for (auto &ev : c->pipe_ev) if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) { g_global_err = "event creation failed"; delete c; return PCB_ERR_CUDA; }
    for (int k = 0; k < pcb_context::N_SIDE; ++k)
        if (cudaStreamCreateWithFlags(&c->sstream[k], cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&c->ev_join[k], cudaEventDisableTiming) != cudaSuccess) {
            g_global_err = "stream creation failed"; delete c; return PCB_ERR_CUDA;
        }
    if (const char *ns = getenv("PCB_STREAMS")) c->n_pipes = std::max(1, std::min(1 + (int)pcb_context::N_SIDE, atoi(ns)));
    for (auto &sl : c->slots) if (cudaEventCreateWithFlags(&sl.staged, cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming) != cudaSuccess) { g_global_err = "event creation failed"; delete c; return PCB_ERR_CUDA; }
    c->grid_trace = pcb_occupancy_grid(0);
    c->grid_scan = pcb_occupancy_grid(1);
    for (int cls = 0; cls < PCF_NCLS; ++cls) {
        c->grid_fscan[cls] = pcf_occupancy_grid(0, cls); c->grid_fp1[cls] = pcf_occupancy_grid(1, cls); c->grid_fp2[cls] = pcf_occupancy_grid(2, cls);
    }
    if (const char *gl = getenv("PCB_GRID_LIMIT")) {       // tests: few CTAs, so that a small input gives every warp many units
        const int lim = std::max(1, atoi(gl));
        c->grid_trace = std::min(c->grid_trace, lim); c->grid_scan = std::min(c->grid_scan, lim);
        for (int cls = 0; cls < PCF_NCLS; ++cls) {
            c->grid_fscan[cls] = std::min(c->grid_fscan[cls], lim); c->grid_fp1[cls] = std::min(c->grid_fp1[cls], lim); c->grid_fp2[cls] = std::min(c->grid_fp2[cls], lim);
        }
    }
    if (const char *mb = getenv("PCB_SCRATCH_MB")) c->scratch_budget = (uint64_t)std::max(1, atoi(mb)) << 18;   // tests: many batches from a small input
    if (c->d_err.reserve(1) != cudaSuccess || cudaHostAlloc((void **)&c->h_totals, sizeof(long long) * pcb_context::N_SLOTS, cudaHostAllocDefault) != cudaSuccess) {
        g_global_err = "cudaMalloc failed"; delete c; return PCB_ERR_NOMEM;
    }
    *out = c;
    return PCB_OK;
}

extern "C" void pcb_destroy(pcb_context *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream); cudaStreamSynchronize(c->cstream); cudaStreamSynchronize(c->ustream); cudaStreamSynchronize(c->pstream);
    c->d_ad_codes.release(); c->d_ad_off.release(); c->d_ascii.release(); c->d_ascii_t.release(); c->d_arena_t.release();
    for (auto &sl : c->slots) {
        sl.d_arena.release(); sl.d_ascii.release(); sl.d_winlen.release(); sl.d_ids.release(); sl.d_out.release();
        sl.d_keys.release(); sl.d_best.release(); sl.d_cnt.release(); sl.d_identcol.release(); sl.d_hits.release(); sl.d_ident.release(); sl.d_total.release(); sl.d_top2.release(); sl.d_trim.release(); sl.d_aux.release(); sl.d_roff.release(); sl.d_rlen.release();
        if (sl.staged) cudaEventDestroy(sl.staged);
        if (sl.done) cudaEventDestroy(sl.done);
    }
    if (c->h_totals) cudaFreeHost(c->h_totals);
    for (auto &pl : c->plans) { pl.release(); if (pl.ready) cudaEventDestroy(pl.ready); }
    c->alt_plan.release();
    c->d_seg_off.release(); c->d_plan_i64.release(); c->d_seg_len.release(); c->d_seg_ad.release(); c->d_plan_i32.release(); c->d_adinfo.release();
    c->d_scratch.release(); c->d_meta.release(); c->d_err.release();
    for (int k = 0; k < pcb_context::N_SIDE; ++k) { c->d_scratch_s[k].release(); c->d_meta_s[k].release(); }
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->kev0); cudaEventDestroy(c->kev1);
    for (auto &ev : c->pipe_ev) if (ev) cudaEventDestroy(ev);
    cudaStreamDestroy(c->cstream);
    cudaStreamDestroy(c->ustream);
    cudaStreamDestroy(c->pstream);
    if (c->hstream) cudaStreamDestroy(c->hstream);
    for (int k = 0; k < pcb_context::N_SIDE; ++k) {
        if (c->sstream[k]) { cudaStreamSynchronize(c->sstream[k]); cudaStreamDestroy(c->sstream[k]); }
        if (c->ev_join[k]) cudaEventDestroy(c->ev_join[k]);
    }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    for (cudaEvent_t ev : c->tasks_done) if (ev) cudaEventDestroy(ev);
    cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" const char *pcb_last_error(pcb_context *c) { return c ? c->err.c_str() : g_global_err.c_str(); }

extern "C" int pcb_select_slot(pcb_context *c, int slot)
{
    if (!c || slot < 0 || slot >= pcb_context::N_SLOTS) return PCB_ERR_ARG;
    c->cur = slot;
    return PCB_OK;
}

extern "C" int pcb_set_scoring(pcb_context *c, int match, int mismatch, int gap_open, int gap_extend)
{
    if (!c) return PCB_ERR_ARG;
    c->sc.match = match; c->sc.mismatch = mismatch; c->sc.gap_open = gap_open; c->sc.gap_extend = gap_extend;
    c->sc.linear = (gap_open == gap_extend);     // seqan/align/global_alignment_unbanded.h:216-219
    c->stamp++;
    return PCB_OK;
}

extern "C" int pcb_set_adapters(pcb_context *c, const char *blob, const int64_t *offsets, int32_t n_adapters)
{
    if (!c || n_adapters < 0 || (n_adapters > 0 && (!blob || !offsets))) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    c->ad_off.assign(offsets, offsets + n_adapters + 1);
    c->n_adapters = n_adapters;
    const int64_t base = n_adapters ? offsets[0] : 0;
    const int64_t total = n_adapters ? offsets[n_adapters] - base : 0;
    for (auto &o : c->ad_off) o -= base;
    for (int i = 0; i < n_adapters; ++i)
        if (c->ad_off[i + 1] < c->ad_off[i]) return c->fail(PCB_ERR_ARG, "adapter offsets must be non-decreasing");
    c->ad_fast.assign((size_t)n_adapters, 0);
    for (int i = 0; i < n_adapters; ++i) {
        const int64_t len = c->ad_off[i + 1] - c->ad_off[i];
        bool ok = len >= 1 && len <= PCF_MAX_N2;
        for (int64_t k = c->ad_off[i]; ok && k < c->ad_off[i + 1]; ++k) ok = strchr("ACGTUacgtu", blob[base + k]) != nullptr && blob[base + k] != 0;
        c->ad_fast[(size_t)i] = (uint8_t)(ok ? pcf_class_of_len((long)len) : PCF_NCLS);
    }
    c->stamp++;
    PCB_CUDA(c, c->d_ad_codes.reserve((size_t)total + 16));
    PCB_CUDA(c, c->d_ad_off.reserve((size_t)n_adapters + 1));
    PCB_CUDA(c, c->d_ascii.reserve((size_t)total + 16));
    if (total) PCB_CUDA(c, cudaMemcpyAsync(c->d_ascii.p, blob + base, (size_t)total, cudaMemcpyHostToDevice, c->stream));
    PCB_CUDA(c, cudaMemcpyAsync(c->d_ad_off.p, c->ad_off.data(), sizeof(int64_t) * (n_adapters + 1), cudaMemcpyHostToDevice, c->stream));
    PCB_CUDA(c, pcb_launch_encode(c->d_ascii.p, c->d_ad_codes.p, (uint64_t)total, c->stream));
    c->cnt.kernel_launches += total ? 1 : 0;
    c->cnt.h2d_bytes += total + (int64_t)sizeof(int64_t) * (n_adapters + 1);
    PCB_CUDA(c, cudaStreamSynchronize(c->stream));     // d_ascii is reused by the staging calls
    return PCB_OK;
}

// ------------------------------------------------------------------------------------------------
// range checks: the kernels compute in signed 16-bit lanes
// ------------------------------------------------------------------------------------------------
static int check_range(pcb_context *c, int nmax, int mmax)
{
    const PcbScores &s = c->sc;
    const int64_t maxabs = std::max(std::max(std::abs((int64_t)s.match), std::abs((int64_t)s.mismatch)),
                                    std::max(std::abs((int64_t)s.gap_open), std::abs((int64_t)s.gap_extend)));
    // with negative gap scores every DP value is bounded by the adapter length alone (a vertical gap from
    // row 0 is always available); otherwise values can grow with the read length too
    const bool neg_gaps = s.gap_open < 0 && s.gap_extend < 0;
    const int64_t bound = maxabs * ((int64_t)nmax + 3 + (neg_gaps ? 0 : mmax));
    if (bound > 8000)
        return c->fail(PCB_ERR_RANGE, "scoring scheme %d,%d,%d,%d with adapter length %d exceeds the 16-bit DP range",
                       s.match, s.mismatch, s.gap_open, s.gap_extend, nmax);
    if (mmax + nmax > 0x3FFF - 2 || mmax > PCB_MAX_TRACE_M)
        return c->fail(PCB_ERR_RANGE, "segment of %d columns is too long for a single trace launch", mmax);
    return PCB_OK;
}

// Scoring schemes the fast kernels (pcb_fast.cu) take: affine with gapOpen < gapExtend < 0, match > 0, and substitution
// scores that stay in [0, 127] after the anti-diagonal bias b = -gapOpen is added twice.
static bool fast_scheme(const PcbScores &s)
{
    const int b = -s.gap_open;
    // b <= 7: the trace checkpoints hold a DP value in 13 bits (64 + b * (52 + 1000) + 52 * match < 8192) and M - H, clamped
    // to gapExtend - gapOpen + 1, in 3
    return s.gap_open < 0 && s.gap_extend < 0 && s.gap_extend > s.gap_open && b <= 7 && s.match > 0 &&
           s.match > s.mismatch && s.mismatch + 2 * b >= 0 && s.match + 2 * b <= 127 &&
           64 + b * (PCF_MAX_N + PCF_MAX_TRACE_M) + PCF_MAX_N * s.match < 8192;
}
// ... and the two-stripe class: 144 register rows, one more iteration than columns
static bool fast_scheme2(const PcbScores &s)
{
    return fast_scheme(s) && 64 + (-s.gap_open) * (144 + PCF_MAX_TRACE_M + 1) + PCF_MAX_N2 * s.match < 8192;
}

// kernel class of adapter a under the context's scheme (PCB_NO_TWO: the two-stripe class goes to the general kernels)
static int class_of(pcb_context *c, int a, bool fast_ok)
{
    int cls = fast_ok ? c->ad_fast[(size_t)a] : PCF_NCLS;
    if (pcf_class_two(cls) && (!fast_scheme2(c->sc) || getenv("PCB_NO_TWO"))) cls = PCF_NCLS;
    return cls;
}

static int ensure_scratch(pcb_context *c, uint64_t words_per_slot, int grid)
{
    const uint64_t need = words_per_slot * (uint64_t)grid * PCB_BLOCK;
    cudaError_t e = c->d_scratch.reserve((size_t)need);
    if (e != cudaSuccess) { cudaGetLastError(); return c->fail(PCB_ERR_NOMEM, "scratch of %.1f MB: %s", need * 4 / 1e6, cudaGetErrorString(e)); }
    return PCB_OK;
}

static int max_adapter_len(pcb_context *c, const int32_t *ids, int n)
{
    int mx = 0;
    for (int i = 0; i < n; ++i) mx = std::max<int>(mx, (int)(c->ad_off[ids[i] + 1] - c->ad_off[ids[i]]));
    return mx;
}

static void fill_launch_common(pcb_context *c, PcbLaunch &L, bool task_arena = false)
{
    memset(&L, 0, sizeof L);
    L.ad.codes = c->d_ad_codes.p; L.ad.off = c->d_ad_off.p; L.ad.n = c->n_adapters;
    L.sc = c->sc;
    L.arena = task_arena ? c->d_arena_t.p : c->S().d_arena.p; L.arena_len = task_arena ? c->arena_t_len : c->S().arena_len;
    L.out = c->S().d_out.p; L.scan_out = nullptr; L.n_out = c->S().n_results;
    L.aux = c->S().want_aux ? c->S().d_aux.p : nullptr;
    L.scratch = c->d_scratch.p; L.err_flag = c->d_err.p;
}

// launch with optional per-kernel event timing
// cls: 0 .. PCF_NCLS-1 = fast kernels (pcb_fast.cu), PCF_NCLS = general kernels (pcb_kernels.cu)
// A fast trace is two kernels per batch of units (pass 1 -> checkpoints + end cells in scratch, then the rounds); a batch is
// as many units as the scratch budget holds.
static int launch_dp(pcb_context *c, PcbLaunch &L, bool trace, int64_t cells, int cls = PCF_NCLS, int pipe = 0 /* 0: main stream, k: side stream k - 1 */)
{
    const int64_t n_units = L.src.n_units;
    if (n_units == 0) return PCB_OK;
    cudaStream_t st = pipe ? c->sstream[pipe - 1] : c->stream;
    DevBuf<uint32_t> &scr = pipe ? c->d_scratch_s[pipe - 1] : c->d_scratch;
    DevBuf<int32_t> &meta = pipe ? c->d_meta_s[pipe - 1] : c->d_meta;
    const bool piped = c->n_pipes > 1 && !c->ktiming && !getenv("PCB_NO_SIDE") && !L.src.units;      // cross products share the budget between their pipelines
    const uint64_t budget = piped ? c->scratch_budget / (uint64_t)c->n_pipes : c->scratch_budget;
    if (c->ktiming) PCB_CUDA(c, cudaEventRecord(c->kev0, c->stream));
    int launches = 1;
    if (cls < PCF_NCLS && trace) {
        // scratch of a batch of units: checkpoints | end cells (4 words per thread) | sorted items | surviving items.
        // (Measured and dropped: batches as a two-stream pipeline, pass 2 of batch i beside pass 1 of batch i + 1 with half
        //  the CTA slots each: 18.5 ms instead of 15.7 ms for 131072 windows x 108 adapters.)
        const uint64_t per_unit = pcf_unit_words_trace(L.mmax, cls), ck_unit = per_unit - 128 - 1024;
        const int W = PCB_BLOCK / 32;
        const int64_t per_batch = (int64_t)std::max<uint64_t>(1, std::min<uint64_t>({(uint64_t)n_units, budget / per_unit, 1ull << 24}));
        cudaError_t e = scr.reserve((size_t)(per_unit * (uint64_t)per_batch));
        if (e != cudaSuccess) { cudaGetLastError(); return c->fail(PCB_ERR_NOMEM, "trace scratch of %.1f MB: %s", per_unit * per_batch * 4 / 1e6, cudaGetErrorString(e)); }
        L.n_major = L.src.units ? c->n_adapters : L.src.n_ids;
        PCB_CUDA(c, meta.reserve((size_t)pcf_meta_words(L.n_major)));
        uint32_t *const sp = scr.p;
        L.meta = meta.p;
        L.scratch = sp;
        L.entries = reinterpret_cast<uint4 *>(sp + ck_unit * (uint64_t)per_batch);
        L.sorted = reinterpret_cast<PcfItem *>(sp + (ck_unit + 128) * (uint64_t)per_batch);
        L.surv = reinterpret_cast<PcfItem *>(sp + (ck_unit + 128 + 512) * (uint64_t)per_batch);
        launches = 0;
        for (int64_t u0 = 0; u0 < n_units; u0 += per_batch) {
            L.unit0 = u0; L.unit1 = std::min(n_units, u0 + per_batch);
            const int64_t ctas = (L.unit1 - L.unit0 + W - 1) / W;
            L.grid = (int)std::min<int64_t>(c->grid_fp1[cls], ctas);
            PCB_CUDA(c, pcf_launch_pass1(L, cls, st));
            PCB_CUDA(c, pcf_launch_pass2(L, cls, (int)std::min<int64_t>(c->grid_fp2[cls], ctas), st));
            launches += pcf_trace_launches(cls);
        }
    } else {
        L.unit0 = 0; L.unit1 = n_units;
        if (cls < PCF_NCLS) PCB_CUDA(c, pcf_launch_scan(L, cls, c->stream));
        else         PCB_CUDA(c, trace ? pcb_launch_trace(L, c->stream) : pcb_launch_scan(L, c->stream));
    }
    if (c->ktiming) {
        PCB_CUDA(c, cudaEventRecord(c->kev1, c->stream));
        PCB_CUDA(c, cudaEventSynchronize(c->kev1));
        float ms = 0; PCB_CUDA(c, cudaEventElapsedTime(&ms, c->kev0, c->kev1));
        c->cnt.dp_kernel_ms += ms;
    }
    c->cnt.kernel_launches += launches; c->cnt.dp_kernel_launches += launches; c->cnt.dp_cells += cells;
    return PCB_OK;
}

static int check_err_flag(pcb_context *c)
{
    int32_t f = 0;
    PCB_CUDA(c, cudaMemcpyAsync(&f, c->d_err.p, sizeof f, cudaMemcpyDeviceToHost, c->stream));
    PCB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (f) return c->fail(PCB_ERR_RANGE, "device reported alignment-window violation (flag %d)", f);
    return PCB_OK;
}

// ------------------------------------------------------------------------------------------------
// windows x adapters cross product
// ------------------------------------------------------------------------------------------------
extern "C" int pcb_stage_windows(pcb_context *c, const char *windows, int64_t n_windows, int32_t stride, const int32_t *win_len)
{
    if (!c || n_windows < 0 || stride <= 0 || (n_windows > 0 && (!windows || !win_len))) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    int mmax = 0; int64_t sumlen = 0;
    for (int64_t w = 0; w < n_windows; ++w) {
        if (win_len[w] < 0 || win_len[w] > stride) return c->fail(PCB_ERR_ARG, "win_len[%lld] = %d outside [0, stride]", (long long)w, win_len[w]);
        mmax = std::max(mmax, win_len[w]); sumlen += win_len[w];
    }
    c->S().sum_winlen = sumlen;
    const uint64_t bytes = (uint64_t)n_windows * stride;
    PCB_CUDA(c, c->S().d_ascii.reserve((size_t)bytes + 16));
    PCB_CUDA(c, c->S().d_arena.reserve((size_t)bytes + 64));
    PCB_CUDA(c, c->S().d_winlen.reserve((size_t)n_windows + 1));
    if (bytes) {
        PCB_CUDA(c, cudaMemcpyAsync(c->S().d_ascii.p, windows, bytes, cudaMemcpyHostToDevice, c->in_stream()));
        PCB_CUDA(c, cudaMemcpyAsync(c->S().d_winlen.p, win_len, sizeof(int32_t) * n_windows, cudaMemcpyHostToDevice, c->in_stream()));
        if (c->deferred) {               // the copies run beside the kernels of earlier calls; the encode waits for them
            PCB_CUDA(c, cudaEventRecord(c->S().staged, c->ustream));
            PCB_CUDA(c, cudaStreamWaitEvent(c->stream, c->S().staged, 0));
        }
        PCB_CUDA(c, pcb_launch_encode(c->S().d_ascii.p, c->S().d_arena.p, bytes, c->stream));
        c->cnt.kernel_launches++;
    }
    // the kernels fetch bases a few bytes past a segment's end: keep valid codes there
    PCB_CUDA(c, cudaMemsetAsync(c->S().d_arena.p + bytes, 0, 64, c->stream));
    c->cnt.h2d_bytes += (int64_t)bytes + (int64_t)sizeof(int32_t) * n_windows;
    c->S().arena_len = bytes; c->S().n_windows = n_windows; c->S().stride = stride; c->S().win_mmax = mmax;
    return PCB_OK;
}

extern "C" int pcb_stage_ids(pcb_context *c, const int32_t *ids, int32_t n_ids)
{
    if (!c || n_ids < 0 || (n_ids > 0 && !ids)) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    for (int i = 0; i < n_ids; ++i)
        if (ids[i] < 0 || ids[i] >= c->n_adapters) return c->fail(PCB_ERR_ARG, "adapter id %d out of range", ids[i]);
    c->S().h_ids.assign(ids, ids + n_ids);
    c->S().n_ids = n_ids;
    c->S().split_stamp = 0;
    return PCB_OK;
}

// Splits the slot's id list by kernel class and uploads [ids of class 0 | class 1 | class 2 | their positions in the same
// order] (position = index in the caller's list).
static int send_pieces(pcb_context *c, int upto);

static int split_ids(pcb_context *c, bool fast_ok)
{
    pcb_context::Slot &S = c->S();
    if (S.split_stamp == c->stamp) return PCB_OK;
    const int n = S.n_ids;
    std::vector<int32_t> id[PCF_NCLS + 1], pos[PCF_NCLS + 1];
    for (int k = 0; k < n; ++k) {
        const int32_t a = S.h_ids[(size_t)k];
        const int cls = class_of(c, a, fast_ok);
        id[cls].push_back(a); pos[cls].push_back(k);
    }
    S.h_split.clear();
    for (int q = 0; q <= PCF_NCLS; ++q) { S.n_cls[q] = (int32_t)id[q].size(); S.h_split.insert(S.h_split.end(), id[q].begin(), id[q].end()); }
    for (int q = 0; q <= PCF_NCLS; ++q) S.h_split.insert(S.h_split.end(), pos[q].begin(), pos[q].end());
    PCB_CUDA(c, S.d_ids.reserve((size_t)2 * n + 1));
    // h_split lives in the slot until the next split, so the asynchronous copy may read it later
    if (n) {
        PCB_CUDA(c, cudaMemcpyAsync(S.d_ids.p, S.h_split.data(), sizeof(int32_t) * 2 * n, cudaMemcpyHostToDevice, c->in_stream()));
        if (c->deferred) {
            PCB_CUDA(c, cudaEventRecord(S.staged, c->ustream));
            PCB_CUDA(c, cudaStreamWaitEvent(c->stream, S.staged, 0));
        }
    }
    c->cnt.h2d_bytes += (int64_t)sizeof(int32_t) * 2 * n;
    S.split_stamp = c->stamp;
    return PCB_OK;
}

// Cross product of the selected slot.  host_out == NULL: results stay in HBM.  Otherwise the biggest launch is cut
// into window ranges and each range's records travel to host_out on the copy stream while the next range computes.
static int run_windows(pcb_context *c, pcb_result *host_out)
{
    pcb_context::Slot &S = c->S();
    const int64_t n_res = S.n_windows * S.n_ids;
    S.n_results = n_res;
    if (n_res == 0) return PCB_OK;
    const int mmax = std::max(1, S.win_mmax);
    const bool fast_ok = fast_scheme(c->sc) && mmax <= PCF_MAX_TRACE_M && !getenv("PCB_NO_FAST");
    int rc = split_ids(c, fast_ok);
    if (rc) return rc;
    PCB_CUDA(c, S.d_out.reserve((size_t)n_res));
    if (!c->deferred) PCB_CUDA(c, cudaMemsetAsync(c->d_err.p, 0, sizeof(int32_t), c->stream));
    // an announced read blob (pcb_upload_reads) travels under this call's kernels: its pieces go out now, behind this
    // call's own copies on the copy engine
    // How much of it: what the copy engine moves while this call's kernels run (about 4.4 T cell updates/s against 50 GB/s:
    // a byte per 88 cells), so that the NEXT window call's own small copies do not wait behind more blob than that.
    // (Half of the blob per call, the first policy, made the second and third window call of a step wait 4 ms each.)
    if ((host_out || c->deferred) && c->uploaded_len >= 0 && c->pending_blob) {
        int64_t sum_n = 0;
        for (int k = 0; k < S.n_ids; ++k) sum_n += c->ad_off[(size_t)S.h_ids[(size_t)k] + 1] - c->ad_off[(size_t)S.h_ids[(size_t)k]];
        const int64_t piece = std::max<int64_t>(1, (c->uploaded_len + c->n_pieces - 1) / c->n_pieces);
        const int64_t now = S.sum_winlen * sum_n / 88;
        const int k = (int)std::max<int64_t>(1, std::min<int64_t>(c->n_pieces, (now + piece - 1) / piece));
        if ((rc = send_pieces(c, c->pieces_sent + k))) return rc;
    }
    const int n = S.n_ids;
    // classes in ascending order of work: the small launches first (their tails then overlap nothing bigger than
    // themselves), the biggest last (it is the one that is pipelined against the copies)
    int order[PCF_NCLS + 1], n_used = 0;
    int64_t work[PCF_NCLS + 1];
    for (int cls = 0; cls <= PCF_NCLS; ++cls) {
        if (S.n_cls[cls] == 0) continue;
        work[cls] = (int64_t)S.n_cls[cls] * (cls < PCF_NCLS ? pcf_class_rows(cls) : 4 * PCB_RT);
        order[n_used++] = cls;
    }
    std::sort(order, order + n_used, [&](int x, int y) { return work[x] < work[y]; });
    // Several pipelines.  A fast class's trace is nine launches that wait for each other (pass 1, sort, rounds, sort, rounds),
    // half of them small: alone they leave SMs idle at every kernel boundary and tail.  So a cross product is dealt to
    // n_pipes streams (own scratch each): the classes in descending order of work, the biggest cut into n_pipes parts by
    // adapters when results stay on the device, each job to the least loaded stream.  Measured on B200 (131,072 windows x
    // the 119 start sequences): 19.3 ms on one stream, 15.6 on two, 15.2 on three or four, 15.3 on eight.  The general
    // class shares the main stream's scratch and stays there.
    struct Job { int cls, start, cnt, pipe; bool last; };
    std::vector<Job> jobs;
    const int n_pipes = (c->ktiming || getenv("PCB_NO_SIDE")) ? 1 : c->n_pipes;
    {
        int64_t load[1 + pcb_context::N_SIDE] = {};
        auto place = [&](int cls, int start, int cnt, bool main_only, bool last) {
            int best = 0;
            if (!main_only) for (int p = 1; p < n_pipes; ++p) if (load[p] < load[best]) best = p;
            load[best] += (int64_t)cnt * (cls < PCF_NCLS ? pcf_class_rows(cls) : 4 * PCB_RT);
            jobs.push_back({cls, start, cnt, best, last});
        };
        for (int oi = n_used - 1; oi >= 0; --oi) {           // biggest first
            const int cls = order[oi], cnt = S.n_cls[cls];
            int start = 0;
            for (int q = 0; q < cls; ++q) start += S.n_cls[q];
            const bool biggest = oi == n_used - 1;
            static const int env_parts = [] { const char *e = getenv("PCB_PARTS"); return e ? atoi(e) : 0; }();
            const int want_parts = env_parts > 0 ? env_parts : n_pipes;
            const int parts = (biggest && n_pipes > 1 && !host_out && cls < PCF_NCLS && !getenv("PCB_NO_SPLIT")) ? std::max(1, std::min(want_parts, cnt / 8)) : 1;
            for (int q = 0; q < parts; ++q) {
                const int a = (int)((int64_t)cnt * q / parts), b = (int)((int64_t)cnt * (q + 1) / parts);
                place(cls, start + a, b - a, cls == PCF_NCLS || (biggest && host_out), biggest && q == parts - 1 && host_out);
            }
        }
        // side jobs are queued first, the main stream's last (its result copies wait for the side streams)
        std::stable_sort(jobs.begin(), jobs.end(), [](const Job &x, const Job &y) { return (x.pipe != 0) > (y.pipe != 0) || ((x.pipe != 0) == (y.pipe != 0) && !x.last && y.last); });
    }
    bool forked = false, joined = false;
    for (const Job &job : jobs) forked |= job.pipe != 0;
    if (forked) {
        PCB_CUDA(c, cudaEventRecord(c->ev_fork, c->stream));
        for (int p = 1; p < n_pipes; ++p) PCB_CUDA(c, cudaStreamWaitEvent(c->sstream[p - 1], c->ev_fork, 0));
    }
    auto join_side = [&]() -> int {                          // the side jobs are all queued: result copies wait for them too
        for (int p = 1; p < n_pipes; ++p) {
            PCB_CUDA(c, cudaEventRecord(c->ev_join[p - 1], c->sstream[p - 1]));
            PCB_CUDA(c, cudaStreamWaitEvent(c->cstream, c->ev_join[p - 1], 0));
        }
        joined = true;
        return PCB_OK;
    };
    bool first_cells = true;
    for (const Job &job : jobs) {
        const int cls = job.cls, cnt = job.cnt, start = job.start;
        if (job.pipe == 0 && forked && !joined && (rc = join_side())) return rc;
        const int32_t *h_list = S.h_split.data() + start;
        const int nmax = max_adapter_len(c, h_list, cnt);
        const int smax = std::max(1, (nmax + PCB_RT - 1) / PCB_RT);
        if (cls == PCF_NCLS && (rc = check_range(c, nmax, mmax))) return rc;
        const int gmax = cls < PCF_NCLS ? c->grid_fp1[cls] : c->grid_trace;
        const int64_t warps = (int64_t)gmax * (PCB_BLOCK / 32);
        int64_t sum_n = 0;                 // algorithmic cell count: sum over windows of len x sum over ids of adapter length
        for (int i = 0; i < cnt; ++i) sum_n += c->ad_off[h_list[i] + 1] - c->ad_off[h_list[i]];
        (void)first_cells;
        // window ranges of this launch (multiples of 64 windows; each range keeps every warp busy for >= 8 units)
        const int64_t groups_all = (S.n_windows + 63) / 64;
        int64_t n_ranges = 1;
        // deferred mode: the copy of a call's results runs under the next call's kernels anyway, and every extra launch costs a tail
        const int max_ranges = c->deferred ? 2 : 8;
        if (host_out && job.last) n_ranges = std::max<int64_t>(1, std::min<int64_t>(max_ranges, groups_all * cnt / (8 * warps)));
        const int64_t groups_per = (groups_all + n_ranges - 1) / n_ranges;
        for (int64_t r = 0; r < n_ranges; ++r) {
            const int64_t w0 = r * groups_per * 64, w1 = std::min<int64_t>(S.n_windows, (r + 1) * groups_per * 64);
            if (w0 >= w1) break;
            const int64_t n_units = ((w1 - w0 + 63) / 64) * cnt;
            const int grid = (int)std::min<int64_t>(gmax, (n_units + PCB_BLOCK / 32 - 1) / (PCB_BLOCK / 32));
            if (cls == PCF_NCLS && (rc = ensure_scratch(c, pcb_scratch_words_trace(mmax, smax), grid))) return rc;
            PcbLaunch L; fill_launch_common(c, L);
            L.arena = S.d_arena.p + (uint64_t)w0 * S.stride;
            L.out = S.d_out.p + w0 * n; L.n_out = S.n_results - w0 * n;
            L.src.units = nullptr; L.src.tasks = nullptr;
            L.src.n_windows = w1 - w0; L.src.stride = S.stride; L.src.win_len = S.d_winlen.p + w0;
            L.src.ids = S.d_ids.p + start; L.src.n_ids = cnt;
            L.src.id_pos = S.d_ids.p + n + start; L.src.n_ids_total = n;
            L.src.n_units = n_units;
            L.ad.max_len = nmax; L.mmax = mmax; L.smax = smax; L.grid = grid;
            if ((rc = launch_dp(c, L, true, r == 0 ? S.sum_winlen * sum_n : 0, cls, job.pipe))) return rc;
            if (host_out && job.last) {
                cudaEvent_t ev = c->pipe_ev[r % 8];
                PCB_CUDA(c, cudaEventRecord(ev, c->stream));
                PCB_CUDA(c, cudaStreamWaitEvent(c->cstream, ev, 0));
                PCB_CUDA(c, cudaMemcpyAsync(host_out + w0 * n, S.d_out.p + w0 * n, sizeof(pcb_result) * (size_t)((w1 - w0) * n),
                                            cudaMemcpyDeviceToHost, c->cstream));
            }
        }
    }
    if (forked && !joined && (rc = join_side())) return rc;
    if (forked) for (int p = 1; p < n_pipes; ++p) PCB_CUDA(c, cudaStreamWaitEvent(c->stream, c->ev_join[p - 1], 0));      // the whole call is complete when the main stream gets here
    if (host_out) {
        c->cnt.d2h_bytes += (int64_t)sizeof(pcb_result) * n_res;
        if (!c->deferred) PCB_CUDA(c, cudaStreamSynchronize(c->cstream));
    }
    return PCB_OK;
}

// Copies the hit lists of the queued pcb_align_windows_hits calls (all of them, or one slot's), now that their sizes are on
// the host.
static int collect_hits(pcb_context *c, int only_slot = -1)
{
    if (c->pending_hits.empty()) return PCB_OK;
    cudaStream_t st = only_slot < 0 ? c->cstream : c->hstream;
    if (only_slot < 0) PCB_CUDA(c, cudaStreamSynchronize(c->cstream));
    int rc = PCB_OK;
    std::vector<pcb_context::PendingHits> rest;
    for (const auto &p : c->pending_hits) {
        if (only_slot >= 0 && p.idx != only_slot) { rest.push_back(p); continue; }
        const int64_t total = (int64_t)c->h_totals[p.idx], n = std::min<int64_t>(total, p.cap);
        *p.n_hits = total;
        if (n > 0) PCB_CUDA(c, cudaMemcpyAsync(p.hits, p.d_hits, sizeof(pcb_hit) * (size_t)n, cudaMemcpyDeviceToHost, st));
        c->cnt.d2h_bytes += (int64_t)sizeof(pcb_hit) * n;
        if (total > p.cap) rc = c->fail(PCB_ERR_RANGE, "%lld hits, room for %lld", (long long)total, (long long)p.cap);
    }
    c->pending_hits.swap(rest);
    PCB_CUDA(c, cudaStreamSynchronize(st));
    return rc;
}

// Deferred mode: waits for ONE slot's queued pcb_align_windows_hits call and delivers its results, while the calls of the
// other slots keep running (the caller can derive that slot's trims and queue its phase C beside them).  Device-side
// errors are still reported by pcb_wait.
extern "C" int pcb_wait_slot(pcb_context *c, int slot)
{
    if (!c || slot < 0 || slot >= pcb_context::N_SLOTS) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    PCB_CUDA(c, cudaEventSynchronize(c->slots[slot].done));
    return collect_hits(c, slot);
}

extern "C" int pcb_run_windows_resident(pcb_context *c)
{
    if (!c) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    return run_windows(c, nullptr);
}

extern "C" int pcb_fetch_results(pcb_context *c, pcb_result *out, int64_t n_results)
{
    if (!c || n_results < 0 || (n_results > 0 && !out)) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    if (n_results > c->S().n_results) return c->fail(PCB_ERR_ARG, "asked for %lld results, last run produced %lld", (long long)n_results, (long long)c->S().n_results);
    if (n_results) PCB_CUDA(c, cudaMemcpyAsync(out, c->S().d_out.p, sizeof(pcb_result) * n_results, cudaMemcpyDeviceToHost, c->stream));
    c->cnt.d2h_bytes += (int64_t)sizeof(pcb_result) * n_results;
    return check_err_flag(c);
}

extern "C" int pcb_align_windows(pcb_context *c, const char *windows, int64_t n_windows, int32_t stride,
                                 const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids, pcb_result *out)
{
    int rc;
    if ((rc = pcb_stage_windows(c, windows, n_windows, stride, win_len))) return rc;
    if ((rc = pcb_stage_ids(c, adapter_ids, n_ids))) return rc;
    if (n_windows * (int64_t)n_ids > 0 && !out) return PCB_ERR_ARG;
    if ((rc = run_windows(c, out))) return rc;
    return c->deferred ? PCB_OK : check_err_flag(c);        // deferred: pcb_wait reports
}

// ------------------------------------------------------------------------------------------------
// the cross product, reduced on the device (pcb_reduce.cu)
// ------------------------------------------------------------------------------------------------
// windows == NULL: the slot's windows are staged already (pcb_stage_windows / pcb_stage_windows_from_reads)
static int stage_or_staged(pcb_context *c, const char *windows, int64_t n_windows, int32_t stride, const int32_t *win_len)
{
    if (windows || !c) return pcb_stage_windows(c, windows, n_windows, stride, win_len);
    if (n_windows != c->S().n_windows) return c->fail(PCB_ERR_ARG, "%lld windows asked for, %lld staged in this slot", (long long)n_windows, (long long)c->S().n_windows);
    return PCB_OK;
}

extern "C" int pcb_align_windows_best(pcb_context *c, const char *windows, int64_t n_windows, int32_t stride,
                                      const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids, pcb_best *best)
{
    int rc;
    if ((rc = stage_or_staged(c, windows, n_windows, stride, win_len))) return rc;
    if ((rc = pcb_stage_ids(c, adapter_ids, n_ids))) return rc;
    if (n_ids > 0 && !best) return PCB_ERR_ARG;
    if ((rc = run_windows(c, nullptr))) return rc;
    if (n_ids == 0) return PCB_OK;
    pcb_context::Slot &S = c->S();
    PCB_CUDA(c, S.d_keys.reserve((size_t)n_ids));
    PCB_CUDA(c, S.d_best.reserve((size_t)n_ids));
    PCB_CUDA(c, pcb_launch_best(S.d_out.p, n_windows, n_ids, S.d_keys.p, S.d_best.p, c->stream));
    c->cnt.kernel_launches += 2;
    PCB_CUDA(c, cudaMemcpyAsync(best, S.d_best.p, sizeof(pcb_best) * n_ids, cudaMemcpyDeviceToHost, c->stream));
    c->cnt.d2h_bytes += (int64_t)sizeof(pcb_best) * n_ids;
    return c->deferred ? PCB_OK : check_err_flag(c);
}

extern "C" int pcb_align_windows_hits(pcb_context *c, const char *windows, int64_t n_windows, int32_t stride,
                                      const int32_t *win_len, const int32_t *adapter_ids, int32_t n_ids,
                                      const pcb_hit_filter *filter, const uint8_t *want_ident,
                                      pcb_hit *hits, int64_t hits_cap, int64_t *n_hits, uint16_t *ident)
{
    int rc;
    if (!c || !filter || !n_hits || hits_cap < 0 || (hits_cap > 0 && !hits)) return PCB_ERR_ARG;
    if ((rc = stage_or_staged(c, windows, n_windows, stride, win_len))) return rc;
    if ((rc = pcb_stage_ids(c, adapter_ids, n_ids))) return rc;
    if ((rc = run_windows(c, nullptr))) return rc;
    pcb_context::Slot &S = c->S();
    *n_hits = 0;
    int32_t n_want = 0;
    S.h_identcol.assign((size_t)std::max(1, n_ids), -1);
    if (want_ident) for (int k = 0; k < n_ids; ++k) if (want_ident[k]) S.h_identcol[(size_t)k] = n_want++;
    S.n_want = n_want;                                      // ident == NULL: the identities stay on the device (pcb_fetch_top2)
    S.hit_filter = *filter; S.has_filter = true;
    if (n_windows == 0 || n_ids == 0) return PCB_OK;
    const int64_t nb = (n_windows + 1023) / 1024;
    PCB_CUDA(c, S.d_cnt.reserve((size_t)(n_windows + nb + 2)));
    PCB_CUDA(c, S.d_identcol.reserve((size_t)n_ids));
    PCB_CUDA(c, S.d_hits.reserve((size_t)std::max<int64_t>(1, hits_cap)));
    PCB_CUDA(c, S.d_ident.reserve((size_t)std::max<int64_t>(1, n_windows * n_want)));
    PCB_CUDA(c, S.d_total.reserve(1));
    PCB_CUDA(c, cudaMemcpyAsync(S.d_identcol.p, S.h_identcol.data(), sizeof(int32_t) * n_ids, cudaMemcpyHostToDevice, c->stream));
    PCB_CUDA(c, pcb_launch_hits(S.d_out.p, n_windows, n_ids, *filter, S.d_cnt.p, S.d_cnt.p + n_windows + 1, S.d_total.p, S.d_hits.p, hits_cap,
                                S.d_identcol.p, n_want, n_want ? S.d_ident.p : nullptr, c->d_err.p, c->stream));
    c->cnt.kernel_launches += 5;
    // results travel on the copy stream, beside the kernels of the next call
    PCB_CUDA(c, cudaEventRecord(S.done, c->stream));
    PCB_CUDA(c, cudaStreamWaitEvent(c->cstream, S.done, 0));
    if (n_want && ident) PCB_CUDA(c, cudaMemcpyAsync(ident, S.d_ident.p, sizeof(uint16_t) * (size_t)(n_windows * n_want), cudaMemcpyDeviceToHost, c->cstream));
    PCB_CUDA(c, cudaMemcpyAsync(c->h_totals + c->cur, S.d_total.p, sizeof(long long), cudaMemcpyDeviceToHost, c->cstream));
    c->cnt.d2h_bytes += (ident ? (int64_t)sizeof(uint16_t) * n_windows * n_want : 0) + 8;
    c->pending_hits.push_back({S.d_hits.p, hits, hits_cap, n_hits, c->cur});
    PCB_CUDA(c, cudaEventRecord(S.done, c->cstream));
    if (c->deferred) return PCB_OK;
    if ((rc = collect_hits(c))) return rc;
    return check_err_flag(c);
}

extern "C" int pcb_fetch_top2(pcb_context *c, pcb_top2 *top2, int64_t n_windows)
{
    if (!c || n_windows < 0 || (n_windows > 0 && !top2)) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    pcb_context::Slot &S = c->S();
    if (n_windows != S.n_windows) return c->fail(PCB_ERR_ARG, "asked for %lld windows, the slot's last call had %lld", (long long)n_windows, (long long)S.n_windows);
    if (n_windows == 0) return PCB_OK;
    PCB_CUDA(c, S.d_top2.reserve((size_t)n_windows));
    PCB_CUDA(c, pcb_launch_top2(S.d_ident.p, n_windows, S.n_want, S.d_top2.p, c->stream));
    c->cnt.kernel_launches++;
    PCB_CUDA(c, cudaMemcpyAsync(top2, S.d_top2.p, sizeof(pcb_top2) * (size_t)n_windows, cudaMemcpyDeviceToHost, c->stream));
    c->cnt.d2h_bytes += (int64_t)sizeof(pcb_top2) * n_windows;
    if (!c->deferred) PCB_CUDA(c, cudaStreamSynchronize(c->stream));
    return PCB_OK;
}

// The trim amounts of the slot's last pcb_align_windows_hits call (same filter), reduced on the device; delivered like that
// call's fixed-size results: at return, or in deferred mode with pcb_wait_slot / pcb_wait.
extern "C" int pcb_fetch_trims(pcb_context *c, int32_t *trim, int64_t n_windows)
{
    if (!c || n_windows < 0 || (n_windows > 0 && !trim)) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    pcb_context::Slot &S = c->S();
    if (!S.has_filter) return c->fail(PCB_ERR_ARG, "pcb_fetch_trims without a pcb_align_windows_hits call in this slot");
    if (n_windows != S.n_windows) return c->fail(PCB_ERR_ARG, "asked for %lld windows, the slot's last call had %lld", (long long)n_windows, (long long)S.n_windows);
    if (n_windows == 0) return PCB_OK;
    PCB_CUDA(c, S.d_trim.reserve((size_t)n_windows));
    if (S.n_ids > 0) PCB_CUDA(c, pcb_launch_trims(S.d_out.p, n_windows, S.n_ids, S.hit_filter, S.d_trim.p, c->stream));
    else PCB_CUDA(c, cudaMemsetAsync(S.d_trim.p, 0xFF, sizeof(int32_t) * (size_t)n_windows, c->stream));
    c->cnt.kernel_launches++;
    PCB_CUDA(c, cudaEventRecord(S.done, c->stream));
    PCB_CUDA(c, cudaStreamWaitEvent(c->cstream, S.done, 0));
    PCB_CUDA(c, cudaMemcpyAsync(trim, S.d_trim.p, sizeof(int32_t) * (size_t)n_windows, cudaMemcpyDeviceToHost, c->cstream));
    c->cnt.d2h_bytes += (int64_t)sizeof(int32_t) * n_windows;
    PCB_CUDA(c, cudaEventRecord(S.done, c->cstream));       // pcb_wait_slot waits for this copy too
    if (!c->deferred) PCB_CUDA(c, cudaStreamSynchronize(c->cstream));
    return PCB_OK;
}

extern "C" int pcb_timer_start(pcb_context *c) { if (!c) return PCB_ERR_ARG; cudaSetDevice(c->device); PCB_CUDA(c, cudaEventRecord(c->ev0, c->stream)); return PCB_OK; }
extern "C" int pcb_timer_stop(pcb_context *c, float *ms)
{
    if (!c || !ms) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    PCB_CUDA(c, cudaEventRecord(c->ev1, c->stream));
    PCB_CUDA(c, cudaEventSynchronize(c->ev1));
    PCB_CUDA(c, cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return PCB_OK;
}
extern "C" int pcb_synchronize(pcb_context *c)
{
    if (!c) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    for (cudaStream_t st : {c->pstream, c->ustream, c->stream, c->cstream}) PCB_CUDA(c, cudaStreamSynchronize(st));
    return PCB_OK;
}
extern "C" int pcb_get_counters(pcb_context *c, pcb_counters *out) { if (!c || !out) return PCB_ERR_ARG; *out = c->cnt; return PCB_OK; }
extern "C" int pcb_reset_counters(pcb_context *c) { if (!c) return PCB_ERR_ARG; memset(&c->cnt, 0, sizeof c->cnt); return PCB_OK; }
extern "C" int pcb_set_kernel_timing(pcb_context *c, int on) { if (!c) return PCB_ERR_ARG; c->ktiming = on != 0; return PCB_OK; }

// ------------------------------------------------------------------------------------------------
// explicit task list (any segment length).
//   len <= direct_max : one trace task.
//   longer            : phase 1 = score-only scan in column chunks (each chunk re-computes a warm-up margin
//                       so that its owned columns are exact), host picks the best end
//                       cell per task in SeqAn's visiting order, phase 2 = trace task on the window that
//                       provably contains the whole path.
// ------------------------------------------------------------------------------------------------
namespace {

struct WindowBounds { int warm; int span; bool ok; };

// Margins for adapter length n under the current scoring scheme (derivation in DESIGN.md section 4).
WindowBounds window_bounds(const PcbScores &s, int n)
{
    WindowBounds b{0, 0, false};
    if (!(s.gap_open < 0 && s.gap_extend < 0)) return b;
    const int64_t smax = std::max(0, std::max(s.match, s.mismatch));
    const int64_t gH = -(int64_t)std::max(s.gap_open, s.gap_extend);          // cheapest horizontal step, > 0
    const int64_t lb = ((int64_t)std::abs(s.gap_open) + std::abs(s.gap_extend)) * (n + 2);
    b.warm = (int)(n + (smax * n + lb + gH - 1) / gH + 2);
    b.span = (int)(n + (smax * n + gH - 1) / gH + 2);
    b.ok = true;
    return b;
}

}  // namespace

// Puts pieces [pieces_sent, upto) of the pending blob on the copy stream: H2D, encode, `ready` event.
static int send_pieces(pcb_context *c, int upto)
{
    if (!c->pending_blob) return PCB_OK;
    for (int p = c->pieces_sent; p < upto && p < c->n_pieces; ++p) {
        pcb_context::TaskPlan &P = c->plans[p];
        if (P.b1 > P.b0) {
            PCB_CUDA(c, cudaMemcpyAsync(c->d_ascii_t.p + P.b0, c->pending_blob + P.b0, (size_t)(P.b1 - P.b0), cudaMemcpyHostToDevice, c->ustream));
        }
        P.encoded = false;            // the encode runs on the main stream, ahead of the piece's kernels (pcb_run_tasks_resident)
        PCB_CUDA(c, cudaEventRecord(P.ready, c->ustream));
        c->pieces_sent = p + 1;
    }
    if (c->pieces_sent >= c->n_pieces) c->pending_blob = nullptr;
    return PCB_OK;
}

// The read blob of the explicit-task path.  n_pieces > 1: it travels in pieces on the upload stream and each piece gets
// its own plan (the tasks whose segment ends inside it), so that the kernels of piece p run while piece p + 1 is still
// on the PCIe bus.  Only piece 0 is sent here; the rest follows from send_pieces (after the plans exist, or pumped by
// the window calls of the same step: small copies issued later would otherwise queue behind the whole blob).
static int setup_blob(pcb_context *c, const char *seq_blob, int64_t blob_len, int n_pieces, int send_now, bool one_plan)
{
    for (auto &pl : c->plans) pl.valid = false;
    PCB_CUDA(c, c->d_ascii_t.reserve((size_t)blob_len + 16));
    PCB_CUDA(c, c->d_arena_t.reserve((size_t)blob_len + 64));
    n_pieces = std::max(1, std::min<int>(n_pieces, pcb_context::MAX_PIECES));
    c->n_pieces = n_pieces;
    c->plan_pieces = one_plan ? 1 : n_pieces;
    c->pending_blob = nullptr;
    PCB_CUDA(c, cudaMemsetAsync(c->d_arena_t.p + blob_len, 0, 64, c->stream));   // valid codes past the last segment
    if (n_pieces == 1) {
        if (blob_len) {
            PCB_CUDA(c, cudaMemcpyAsync(c->d_ascii_t.p, seq_blob, (size_t)blob_len, cudaMemcpyHostToDevice, c->stream));
            PCB_CUDA(c, pcb_launch_encode(c->d_ascii_t.p, c->d_arena_t.p, (uint64_t)blob_len, c->stream));
            c->cnt.kernel_launches++;
        }
        c->plans[0].b0 = 0; c->plans[0].b1 = blob_len;
        c->plans[0].end_lo = -1; c->plans[0].end_hi = blob_len; c->plans[0].has_ready = false; c->plans[0].encoded = true;
    } else {
        // the upload stream starts after whatever the main stream still does with the arena
        PCB_CUDA(c, cudaEventRecord(c->pipe_ev[0], c->stream));
        PCB_CUDA(c, cudaStreamWaitEvent(c->ustream, c->pipe_ev[0], 0));
        const int64_t per = ((blob_len + n_pieces - 1) / n_pieces + 15) & ~(int64_t)15;
        for (int p = 0; p < n_pieces; ++p) {
            pcb_context::TaskPlan &P = c->plans[p];
            P.b0 = std::min<int64_t>(blob_len, p * per); P.b1 = p == n_pieces - 1 ? blob_len : std::min<int64_t>(blob_len, (p + 1) * per);
            if (!P.ready) PCB_CUDA(c, cudaEventCreateWithFlags(&P.ready, cudaEventDisableTiming));
            P.has_ready = true;
            // the kernels fetch bases up to 8 bytes past a segment's end: a task waits for the piece that holds those too
            P.end_lo = p == 0 ? -1 : P.b0 - 16; P.end_hi = p == n_pieces - 1 ? blob_len : P.b1 - 16;
            if (one_plan) { P.end_lo = p == 0 ? -1 : blob_len; P.end_hi = blob_len; }      // plan 0 takes every task
        }
        c->pending_blob = seq_blob; c->pieces_sent = 0;
    }
    c->cnt.h2d_bytes += blob_len;
    c->arena_t_len = (uint64_t)blob_len;
    return send_pieces(c, send_now);
}

// The task list over the blob that setup_blob put (or is putting) into the arena.
static int stage_segments(pcb_context *c, int64_t blob_len, const int64_t *seg_off, const int32_t *seg_len,
                          const int32_t *task_adapter, int64_t n_tasks)
{
    for (auto &pl : c->plans) pl.valid = false;
    c->alt_plan.valid = false;
    c->n_tasks = 0;
    // result slots, task and chunk indices are 32-bit on the device (PcbTask.out_idx, PcbLong.chunk0 / win_slot)
    if (n_tasks >= (1ll << 31) - 64) return c->fail(PCB_ERR_RANGE, "%lld tasks in one call: the task path indexes with 32 bits", (long long)n_tasks);
    PCB_CUDA(c, c->d_seg_off.reserve((size_t)n_tasks + 1));
    PCB_CUDA(c, c->d_seg_len.reserve((size_t)n_tasks + 1));
    PCB_CUDA(c, c->d_seg_ad.reserve((size_t)n_tasks + 1));
    // deferred mode: the copies and the planner run on their own stream, beside the kernels of earlier calls
    // (the plan set alternates between the task calls on an announced blob, so the planner of this call only waits for
    //  the kernels of the call before the last one)
    if (c->deferred && c->plan_pieces == 1 && c->uploaded_len >= 0) {
        c->plan_flip ^= 1;
        c->alt_plan.b0 = c->plans[0].b0; c->alt_plan.b1 = c->plans[0].b1;
        c->alt_plan.end_lo = c->plans[0].end_lo; c->alt_plan.end_hi = c->plans[0].end_hi;
        PCB_CUDA(c, cudaStreamWaitEvent(c->pstream, c->tasks_done[c->plan_flip], 0));
    } else {
        c->plan_flip = 0;
        if (c->deferred) for (cudaEvent_t ev : c->tasks_done) PCB_CUDA(c, cudaStreamWaitEvent(c->pstream, ev, 0));
    }
    if (n_tasks) {          // asynchronous from pinned memory: the host-side checks below run under the copies
        PCB_CUDA(c, cudaMemcpyAsync(c->d_seg_off.p, seg_off, sizeof(int64_t) * n_tasks, cudaMemcpyHostToDevice, c->plan_stream()));
        PCB_CUDA(c, cudaMemcpyAsync(c->d_seg_len.p, seg_len, sizeof(int32_t) * n_tasks, cudaMemcpyHostToDevice, c->plan_stream()));
        PCB_CUDA(c, cudaMemcpyAsync(c->d_seg_ad.p, task_adapter, sizeof(int32_t) * n_tasks, cudaMemcpyHostToDevice, c->plan_stream()));
    }
    int64_t long_cols = 0, cells = 0, max_end = 0;
    for (int64_t t = 0; t < n_tasks; ++t) {
        const bool bad_seg = seg_len[t] < 0 || seg_off[t] < 0 || seg_off[t] + seg_len[t] > blob_len;
        const bool bad_ad = task_adapter[t] < 0 || task_adapter[t] >= c->n_adapters;
        if (bad_seg || bad_ad) {
            cudaStreamSynchronize(c->plan_stream());               // the caller's buffers are still being read
            if (bad_seg) return c->fail(PCB_ERR_ARG, "task %lld: segment [%lld, +%d) outside the blob", (long long)t, (long long)seg_off[t], seg_len[t]);
            return c->fail(PCB_ERR_ARG, "task %lld: adapter id %d out of range", (long long)t, task_adapter[t]);
        }
        if (seg_len[t] > 960) long_cols += seg_len[t];
        max_end = std::max<int64_t>(max_end, seg_off[t] + seg_len[t]);
        cells += (int64_t)seg_len[t] * (c->ad_off[(size_t)task_adapter[t] + 1] - c->ad_off[(size_t)task_adapter[t]]);
    }
    c->n_tasks = n_tasks; c->task_long_cols = long_cols; c->task_cells = cells; c->task_max_end = max_end;
    c->cnt.h2d_bytes += 16 * n_tasks;
    return PCB_OK;
}

static int blob_pieces(int64_t blob_len)
{
    const char *pm = getenv("PCB_PIECE_MIN_BYTES");           // tests lower it to reach the pieced path with small inputs
    const int64_t piece_min = pm ? atoll(pm) : (64ll << 20);
    const char *pp = getenv("PCB_PLAN_PIECES");
    // about 128 MB per piece (2.4 ms on the bus): measured on 3.1 GB of ultra-long reads, 4 / 8 / 16 pieces = 69 / 65 / 63 ms per step
    const int by_size = (int)std::max<int64_t>(2, std::min<int64_t>(pcb_context::MAX_PIECES, blob_len >> 27));
    const int pieces = pp ? std::max(1, std::min<int>(atoi(pp), pcb_context::MAX_PIECES)) : by_size;
    return blob_len >= piece_min && blob_len >= 4096 ? pieces : 1;
}

static int stage_tasks(pcb_context *c, const char *seq_blob, int64_t blob_len, const int64_t *seg_off,
                       const int32_t *seg_len, const int32_t *task_adapter, int64_t n_tasks, int n_pieces)
{
    if (!c || blob_len < 0 || n_tasks < 0 || (n_tasks > 0 && (!seg_off || !seg_len || !task_adapter)) || (blob_len > 0 && !seq_blob))
        return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    c->uploaded_len = -1;
    int rc = setup_blob(c, seq_blob, blob_len, n_pieces, 1, false);
    if (rc) return rc;
    return stage_segments(c, blob_len, seg_off, seg_len, task_adapter, n_tasks);
}

extern "C" int pcb_upload_reads(pcb_context *c, const char *seq_blob, int64_t blob_len)
{
    if (!c || blob_len < 0 || (blob_len > 0 && !seq_blob)) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    c->n_tasks = 0;
    // nothing is sent yet: the window calls that follow pump the pieces AFTER their own (small) copies, so that those
    // do not wait behind the blob on the copy engine; pcb_align_tasks sends whatever is left
    int rc = setup_blob(c, seq_blob, blob_len, blob_pieces(blob_len) > 1 ? (int)pcb_context::MAX_PIECES : 1, 0, true);
    if (rc) return rc;
    c->uploaded_len = blob_len;
    return PCB_OK;
}

// The announced blob, all of it, on the device and encoded (main stream order).
static int blob_resident(pcb_context *c)
{
    int rc = send_pieces(c, c->n_pieces);
    if (rc) return rc;
    for (int p = 0; p < c->n_pieces; ++p) {
        pcb_context::TaskPlan &P = c->plans[p];
        if (P.has_ready) PCB_CUDA(c, cudaStreamWaitEvent(c->stream, P.ready, 0));
        if (!P.encoded) {
            if (P.b1 > P.b0) {
                PCB_CUDA(c, pcb_launch_encode(c->d_ascii_t.p + P.b0, c->d_arena_t.p + P.b0, (uint64_t)(P.b1 - P.b0), c->stream));
                c->cnt.kernel_launches++;
            }
            P.encoded = true;
        }
    }
    return PCB_OK;
}

extern "C" int pcb_stage_windows_from_reads(pcb_context *c, const int64_t *read_off, const int32_t *read_len, int64_t n_reads,
                                            int32_t end_size, int32_t at_end)
{
    if (!c || n_reads < 0 || end_size <= 0 || (n_reads > 0 && (!read_off || !read_len))) return PCB_ERR_ARG;
    if (c->uploaded_len < 0) return c->fail(PCB_ERR_ARG, "pcb_stage_windows_from_reads needs the read blob (pcb_upload_reads)");
    cudaSetDevice(c->device);
    pcb_context::Slot &S = c->S();
    const int32_t stride = (end_size + 15) / 16 * 16;
    int mmax = 0; int64_t sumlen = 0;
    for (int64_t r = 0; r < n_reads; ++r) {
        if (read_len[r] < 0 || read_off[r] < 0 || read_off[r] + read_len[r] > c->uploaded_len)
            return c->fail(PCB_ERR_ARG, "read %lld: [%lld, +%d) outside the blob", (long long)r, (long long)read_off[r], read_len[r]);
        const int wl = std::min(read_len[r], end_size);
        mmax = std::max(mmax, wl); sumlen += wl;
    }
    int rc = blob_resident(c);
    if (rc) return rc;
    const uint64_t bytes = (uint64_t)n_reads * stride;
    PCB_CUDA(c, S.d_arena.reserve((size_t)bytes + 64));
    PCB_CUDA(c, S.d_winlen.reserve((size_t)n_reads + 1));
    PCB_CUDA(c, S.d_roff.reserve((size_t)n_reads + 1));
    PCB_CUDA(c, S.d_rlen.reserve((size_t)n_reads + 1));
    if (n_reads) {
        PCB_CUDA(c, cudaMemcpyAsync(S.d_roff.p, read_off, sizeof(int64_t) * n_reads, cudaMemcpyHostToDevice, c->stream));
        PCB_CUDA(c, cudaMemcpyAsync(S.d_rlen.p, read_len, sizeof(int32_t) * n_reads, cudaMemcpyHostToDevice, c->stream));
        PCB_CUDA(c, pcb_launch_gather_windows(c->d_arena_t.p, S.d_roff.p, S.d_rlen.p, n_reads, end_size, at_end, stride, S.d_arena.p, S.d_winlen.p, c->stream));
        c->cnt.kernel_launches++;
    }
    PCB_CUDA(c, cudaMemsetAsync(S.d_arena.p + bytes, 0, 64, c->stream));
    c->cnt.h2d_bytes += 12 * n_reads;
    S.arena_len = bytes; S.n_windows = n_reads; S.stride = stride; S.win_mmax = mmax; S.sum_winlen = sumlen;
    if (!c->deferred) PCB_CUDA(c, cudaStreamSynchronize(c->stream));       // the caller's arrays may go away
    return PCB_OK;
}

extern "C" int pcb_stage_tasks(pcb_context *c, const char *seq_blob, int64_t blob_len, const int64_t *seg_off,
                               const int32_t *seg_len, const int32_t *task_adapter, int64_t n_tasks)
{
    return stage_tasks(c, seq_blob, blob_len, seg_off, seg_len, task_adapter, n_tasks, 1);
}

// ------------------------------------------------------------------------------------------------
// The plan of a staged task list: every launch the run needs, with its task / unit arrays already in HBM.  It depends
// only on the staged segments, the adapter set and the scoring scheme, so it is built once per staging (or when one
// of those changes) and a run is nothing but a sequence of kernel launches on the context's stream: direct traces,
// chunk scans, one merge kernel (best end cell per long read -> its trace-window task), window traces.  No host
// round trip inside a run.
// ------------------------------------------------------------------------------------------------
static int build_task_plan(pcb_context *c, pcb_context::TaskPlan &P, bool first_piece)
{
    P.launches.clear(); P.n_long = 0; P.n_chunks = 0; P.win_task0 = 0; P.scratch_words = 0; P.valid = false;
    const int64_t nt = c->n_tasks;
    if (nt == 0) { P.valid = true; P.stamp = c->stamp; return PCB_OK; }
    const bool fast_sc = fast_scheme(c->sc) && !getenv("PCB_NO_FAST");
    const int A = c->n_adapters;

    // per adapter: length, class, margins
    std::vector<PcbAdInfo> ad((size_t)A);
    const int direct_default = 960;
    for (int a = 0; a < A; ++a) {
        PcbAdInfo &x = ad[(size_t)a];
        x.n = (int)(c->ad_off[(size_t)a + 1] - c->ad_off[(size_t)a]);
        x.cls = class_of(c, a, fast_sc);
        const WindowBounds wb = window_bounds(c->sc, x.n);
        x.warm = wb.warm; x.span = wb.span; x.chunked = wb.ok ? 1 : 0;
        x.direct_max = std::max(direct_default, wb.ok ? wb.warm + wb.span + 8 : 0);
        // what the fast kernels take: direct tasks of <= PCF_MAX_TRACE_M columns (longer segments are scanned in chunks and
        // traced on a window; a window wider than that goes to the general kernel, cls_of below) and scan chunks of
        // <= PCF_MAX_SCAN_M columns including their warm-up margin
        if (x.cls < PCF_NCLS && wb.ok && wb.warm + 4100 > PCF_MAX_SCAN_M) x.cls = PCF_NCLS;
        if (x.cls < PCF_NCLS) x.direct_max = std::min(x.direct_max, PCF_MAX_TRACE_M);
    }
    // chunk length: enough chunks to fill the machine, but long enough to amortise the warm-up margin
    const int chunk = (int)std::min<int64_t>(4096, std::max<int64_t>(1024, c->task_long_cols / (2 * 64 * (int64_t)std::max(1, c->grid_fscan[0]) * (PCB_BLOCK / 32))));

    // pass 1 on the device: items per (kind, adapter, length bucket), widest item per (kind, class)
    const size_t nb = (size_t)PCB_KINDS * A * PCB_LB, nm = (size_t)PCB_KINDS * (PCF_NCLS + 1);
    PCB_CUDA(c, c->d_plan_i32.reserve(2 * nb + nm));
    PCB_CUDA(c, c->d_plan_i64.reserve(nb));
    PCB_CUDA(c, c->d_adinfo.reserve((size_t)A));
    PCB_CUDA(c, cudaMemsetAsync(c->d_plan_i32.p, 0, sizeof(int32_t) * (2 * nb + nm), c->plan_stream()));
    PCB_CUDA(c, cudaMemcpyAsync(c->d_adinfo.p, ad.data(), sizeof(PcbAdInfo) * A, cudaMemcpyHostToDevice, c->plan_stream()));
    PcbPlanArgs pa{};
    pa.seg_off = c->d_seg_off.p; pa.seg_len = c->d_seg_len.p; pa.seg_ad = c->d_seg_ad.p; pa.n_tasks = nt;
    pa.ad = c->d_adinfo.p; pa.n_adapters = A; pa.chunk = chunk; pa.end_lo = P.end_lo; pa.end_hi = P.end_hi;
    pa.count = c->d_plan_i32.p; pa.cursor = c->d_plan_i32.p + nb; pa.mmax = c->d_plan_i32.p + 2 * nb; pa.base = c->d_plan_i64.p;
    PCB_CUDA(c, pcb_launch_plan(pa, false, c->plan_stream()));
    std::vector<int32_t> count(nb), mmax(nm);
    PCB_CUDA(c, cudaMemcpyAsync(count.data(), pa.count, sizeof(int32_t) * nb, cudaMemcpyDeviceToHost, c->plan_stream()));
    PCB_CUDA(c, cudaMemcpyAsync(mmax.data(), pa.mmax, sizeof(int32_t) * nm, cudaMemcpyDeviceToHost, c->plan_stream()));
    PCB_CUDA(c, cudaStreamSynchronize(c->plan_stream()));
    c->cnt.kernel_launches += 2;

    // layout: kind by kind, class by class (the launch order), adapter by adapter, bucket by bucket
    std::vector<int64_t> base(nb, 0);
    std::vector<PcbUnit> all_units;
    int64_t run = 0;
    int rc;
    for (int kind = 0; kind < PCB_KINDS; ++kind) {
        if (kind == 1) pa.scan_task0 = run;
        if (kind == 2) pa.win_task0 = run;
        const bool trace = kind != 1;
        for (int q = PCF_NCLS; q >= 0; --q) {
            // class of an adapter's items of this kind: windows wider than the fast trace kernel takes go to the general one
            auto cls_of = [&](int a) { const PcbAdInfo &x = ad[(size_t)a]; return (kind == 2 && x.warm + x.span > PCF_MAX_TRACE_M) ? PCF_NCLS : x.cls; };
            int64_t total = 0; int nmax = 1, wmax = 1;
            for (int a = 0; a < A; ++a) {
                if (cls_of(a) != q) continue;
                int64_t na = 0;
                for (int lb = 0; lb < PCB_LB; ++lb) na += count[((size_t)kind * A + a) * PCB_LB + lb];
                if (na) { nmax = std::max(nmax, ad[(size_t)a].n); wmax = std::max(wmax, ad[(size_t)a].warm + ad[(size_t)a].span); }
                total += na;
            }
            if (total == 0) continue;
            const bool fast = q < PCF_NCLS;
            const int lmax = kind == 2 ? wmax : std::max(1, kind == 0 ? mmax[(size_t)q] : mmax[(size_t)(PCF_NCLS + 1) + q]);
            if (!fast && (rc = check_range(c, nmax, lmax))) return rc;
            if (fast && lmax > (trace ? PCF_MAX_TRACE_M : PCF_MAX_SCAN_M)) return c->fail(PCB_ERR_RANGE, "internal: segment of %d columns sent to the fast kernels", lmax);
            const int gmax = fast ? (trace ? c->grid_fp1[q] : c->grid_fscan[q]) : (trace ? c->grid_trace : c->grid_scan);
            // fewer tasks per warp when there is not enough work to fill the machine (a warp's time does not depend on its lanes)
            const int64_t warps = (int64_t)gmax * (PCB_BLOCK / 32);
            const int per_unit = (int)std::min<int64_t>(64, std::max<int64_t>(2, (total + warps - 1) / warps));
            pcb_context::TaskPlan::Launch pl{};
            pl.trace = trace; pl.cls = q; pl.kind = kind; pl.cells = 0;
            pl.task0 = run; pl.unit0 = (int64_t)all_units.size();
            for (int a = 0; a < A; ++a) {
                if (cls_of(a) != q) continue;
                int64_t na = 0;
                for (int lb = 0; lb < PCB_LB; ++lb) {
                    const size_t b = ((size_t)kind * A + a) * PCB_LB + lb;
                    base[b] = run + na; na += count[b];
                }
                for (int64_t u = 0; u < na; u += per_unit) {
                    PcbUnit un; un.adapter = a; un.ntask = (int32_t)std::min<int64_t>(per_unit, na - u); un.task0 = run - pl.task0 + u;
                    all_units.push_back(un);
                }
                run += na;
            }
            pl.n_units = (int64_t)all_units.size() - pl.unit0;
            pl.mmax = lmax; pl.nmax = nmax; pl.smax = std::max(1, (nmax + PCB_RT - 1) / PCB_RT);
            pl.grid = (int)std::min<int64_t>(gmax, (pl.n_units + PCB_BLOCK / 32 - 1) / (PCB_BLOCK / 32));
            // (the fast trace sizes its own scratch, batch by batch: launch_dp)
            const uint64_t words = fast ? 1 : (trace ? pcb_scratch_words_trace(lmax, pl.smax) : pcb_scratch_words_scan(lmax, pl.smax));
            P.scratch_words = std::max<uint64_t>(P.scratch_words, words * (uint64_t)pl.grid * PCB_BLOCK);
            P.launches.push_back(pl);
        }
    }
    if (!P.launches.empty() && first_piece) P.launches[0].cells = c->task_cells;      // algorithmic cell updates of the whole run
    if (run >= (1ll << 31) - 64) return c->fail(PCB_ERR_RANGE, "%lld kernel tasks (direct + scan chunks + windows) in one plan: the task path indexes with 32 bits", (long long)run);
    P.n_chunks = pa.win_task0 - pa.scan_task0;
    P.n_long = run - pa.win_task0;
    P.win_task0 = pa.win_task0;

    // pass 2 on the device: write the tasks into their buckets
    PCB_CUDA(c, P.d_tasks.reserve((size_t)run + 1));
    PCB_CUDA(c, P.d_units.reserve(all_units.size() + 1));
    PCB_CUDA(c, P.d_longs.reserve((size_t)P.n_long + 1));
    PCB_CUDA(c, P.d_scan.reserve((size_t)P.n_chunks + 1));
    PCB_CUDA(c, cudaMemcpyAsync(c->d_plan_i64.p, base.data(), sizeof(int64_t) * nb, cudaMemcpyHostToDevice, c->plan_stream()));
    PCB_CUDA(c, cudaMemcpyAsync(P.d_units.p, all_units.data(), sizeof(PcbUnit) * all_units.size(), cudaMemcpyHostToDevice, c->plan_stream()));
    pa.tasks = P.d_tasks.p; pa.longs = P.d_longs.p;
    PCB_CUDA(c, pcb_launch_plan(pa, true, c->plan_stream()));
    c->cnt.h2d_bytes += (int64_t)(sizeof(int64_t) * nb + sizeof(PcbUnit) * all_units.size() + sizeof(PcbAdInfo) * A);
    c->cnt.d2h_bytes += (int64_t)(sizeof(int32_t) * (nb + nm));
    cudaError_t e = c->d_scratch.reserve((size_t)P.scratch_words);
    if (e != cudaSuccess) { cudaGetLastError(); return c->fail(PCB_ERR_NOMEM, "scratch of %.1f MB: %s", P.scratch_words * 4 / 1e6, cudaGetErrorString(e)); }
    PCB_CUDA(c, cudaStreamSynchronize(c->plan_stream()));       // the host vectors go out of scope
    P.valid = true; P.stamp = c->stamp;
    return PCB_OK;
}

extern "C" int pcb_run_tasks_resident(pcb_context *c)
{
    if (!c) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    const size_t nt = (size_t)c->n_tasks;
    c->S().n_results = (int64_t)nt;
    if (nt == 0) return PCB_OK;
    int rc;
    auto plan_of = [&](int p) -> pcb_context::TaskPlan & { return c->plan_pieces == 1 ? c->one_plan() : c->plans[p]; };
    for (int p = 0; p < c->plan_pieces; ++p)
        if ((!plan_of(p).valid || plan_of(p).stamp != c->stamp) && (rc = build_task_plan(c, plan_of(p), p == 0))) return rc;
    if ((rc = send_pieces(c, c->n_pieces))) return rc;
    PCB_CUDA(c, c->S().d_out.reserve(nt));
    uint64_t scratch = 0;
    for (int p = 0; p < c->plan_pieces; ++p) scratch = std::max(scratch, plan_of(p).scratch_words);
    PCB_CUDA(c, c->d_scratch.reserve((size_t)scratch));
    if (!c->deferred) PCB_CUDA(c, cudaMemsetAsync(c->d_err.p, 0, sizeof(int32_t), c->stream));
    auto arrive = [&](pcb_context::TaskPlan &P) -> int {   // this piece of the blob is on the device, as codes
        if (P.has_ready) PCB_CUDA(c, cudaStreamWaitEvent(c->stream, P.ready, 0));
        if (!P.encoded) {
            if (P.b1 > P.b0) {
                PCB_CUDA(c, pcb_launch_encode(c->d_ascii_t.p + P.b0, c->d_arena_t.p + P.b0, (uint64_t)(P.b1 - P.b0), c->stream));
                c->cnt.kernel_launches++;
            }
            P.encoded = true;
        }
        return PCB_OK;
    };
    // one plan over an announced blob: it waits for the pieces its segments reach into, not for the whole blob (a batch that
    // goes through in parts starts the first part's tasks while the later reads are still on the bus; the kernels fetch
    // bases up to 8 bytes past a segment's end)
    if (c->plan_pieces == 1) for (int p = 1; p < c->n_pieces; ++p) if (c->plans[p].b0 < c->task_max_end + 16 && (rc = arrive(c->plans[p]))) return rc;
    for (int p = 0; p < c->plan_pieces; ++p) {
        pcb_context::TaskPlan &P = plan_of(p);
        if ((rc = arrive(c->plans[p]))) return rc;
        bool merged = false;
        for (const auto &pl : P.launches) {
            if (pl.kind == 2 && !merged) {
                PCB_CUDA(c, pcb_launch_merge(P.d_longs.p, P.n_long, P.d_scan.p, P.d_tasks.p + P.win_task0, c->d_err.p, c->stream));
                c->cnt.kernel_launches++;
                merged = true;
            }
            PcbLaunch L; fill_launch_common(c, L, true);
            L.scan_out = P.d_scan.p;
            L.src.units = P.d_units.p + pl.unit0; L.src.tasks = P.d_tasks.p + pl.task0; L.src.n_units = pl.n_units;
            L.ad.max_len = pl.nmax; L.mmax = pl.mmax; L.smax = pl.smax; L.grid = pl.grid;
            if ((rc = launch_dp(c, L, pl.trace, pl.cells, pl.cls))) return rc;
        }
    }
    PCB_CUDA(c, cudaEventRecord(c->tasks_done[c->plan_pieces == 1 ? c->plan_flip : 0], c->stream));
    return PCB_OK;
}

extern "C" int pcb_align_tasks(pcb_context *c, const char *seq_blob, int64_t blob_len, const int64_t *seg_off,
                               const int32_t *seg_len, const int32_t *task_adapter, int64_t n_tasks, pcb_result *out)
{
    int rc;
    if (!seq_blob && c && c->uploaded_len >= 0) {
        // the blob was announced with pcb_upload_reads and is in the arena, or on its way there
        if (blob_len != c->uploaded_len) return c->fail(PCB_ERR_ARG, "blob_len %lld differs from the uploaded blob (%lld)", (long long)blob_len, (long long)c->uploaded_len);
        if (n_tasks < 0 || (n_tasks > 0 && (!seg_off || !seg_len || !task_adapter))) return PCB_ERR_ARG;
        cudaSetDevice(c->device);
        if ((rc = stage_segments(c, blob_len, seg_off, seg_len, task_adapter, n_tasks))) return rc;
    } else {
        // big blobs travel in pieces, each piece's kernels running under the next piece's copy
        if ((rc = stage_tasks(c, seq_blob, blob_len, seg_off, seg_len, task_adapter, n_tasks, blob_pieces(blob_len)))) return rc;
    }
    if ((rc = pcb_run_tasks_resident(c))) return rc;
    if (!c->deferred) return pcb_fetch_results(c, out, n_tasks);
    if (n_tasks > 0 && !out) return PCB_ERR_ARG;
    if (n_tasks) PCB_CUDA(c, cudaMemcpyAsync(out, c->S().d_out.p, sizeof(pcb_result) * n_tasks, cudaMemcpyDeviceToHost, c->stream));
    c->cnt.d2h_bytes += (int64_t)sizeof(pcb_result) * n_tasks;
    return PCB_OK;                      // pcb_wait reports
}

extern "C" int pcb_set_deferred(pcb_context *c, int on)
{
    if (!c) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    for (cudaStream_t st : {c->stream, c->cstream, c->ustream, c->pstream}) PCB_CUDA(c, cudaStreamSynchronize(st));
    PCB_CUDA(c, cudaMemsetAsync(c->d_err.p, 0, sizeof(int32_t), c->stream));
    PCB_CUDA(c, cudaStreamSynchronize(c->stream));
    c->deferred = on != 0;
    return PCB_OK;
}

extern "C" int pcb_wait(pcb_context *c)
{
    if (!c) return PCB_ERR_ARG;
    cudaSetDevice(c->device);
    for (cudaStream_t st : {c->pstream, c->ustream, c->stream, c->cstream}) PCB_CUDA(c, cudaStreamSynchronize(st));
    const int rc_hits = collect_hits(c);
    int32_t f = 0;
    PCB_CUDA(c, cudaMemcpyAsync(&f, c->d_err.p, sizeof f, cudaMemcpyDeviceToHost, c->stream));
    PCB_CUDA(c, cudaMemsetAsync(c->d_err.p, 0, sizeof(int32_t), c->stream));
    PCB_CUDA(c, cudaStreamSynchronize(c->stream));
    if (f) return c->fail(PCB_ERR_RANGE, "device reported alignment-window violation (flag %d)", f);
    return rc_hits;
}

// ------------------------------------------------------------------------------------------------
// result formatting and the reference's two symbols
// ------------------------------------------------------------------------------------------------
extern "C" void pcb_format_result(const pcb_result *r, char *buf, uint64_t buflen)
{
    // ScoredAlignment::getString (alignment.cpp:112-120): std::to_string = "%d" / "%f";
    // identities exactly as alignment.cpp:81,89: 100.0 * count / length, evaluated in double.
    double aligned = 0.0, full = 0.0;
    if (r->read_start != -1) {
        aligned = 100.0 * r->matches / r->aligned_len;
        full = 100.0 * r->matches / r->full_len;
    }
    snprintf(buf, (size_t)buflen, "%d,%d,%d,%d,%d,%f,%f", r->read_start, r->read_end, r->adapter_start,
             r->adapter_end, r->raw_score, aligned, full);
}

extern "C" int pcb_stats_from_path(const char *ops, int64_t n_ops, int64_t head_h, int64_t head_v, int64_t tail_h,
                                   int64_t tail_v, const uint8_t *match_flags, int32_t raw_score, pcb_result *out)
{
    if (!out || n_ops < 0 || (n_ops > 0 && (!ops || !match_flags))) return PCB_ERR_ARG;
    PathAcc acc; path_acc_init(acc);
    int64_t m = head_h + tail_h, n = head_v + tail_v;
    int ei = 0, ej = 0;                                      // core path size: rows / columns it covers
    for (int64_t k = 0; k < n_ops; ++k) { if (ops[k] != 'H') ++ei; if (ops[k] != 'V') ++ej; }
    WalkAcc wa; walk_acc_init(wa);
    uint32_t pos2 = (uint32_t)(2 * ej + 1) | ((uint32_t)(2 * ei + 1) << 16);
    int matches = 0;
    for (int64_t k = n_ops - 1; k >= 0; --k) {              // the kernels walk the path back to front
        const int op = ops[k] == 'D' ? 0 : (ops[k] == 'V' ? 1 : 2);
        path_acc_step(acc, op, match_flags[k]);
        matches += op == 0 && match_flags[k];
        // a gap op extends when the op before it (in path order) is of the same kind
        const uint32_t ext = op != 0 && k > 0 && ops[k - 1] == ops[k];
        walk_acc_step(wa, pos2, op == 0, op == 1, op == 2, ext);
        if (op != 2) ++n;
        if (op != 1) ++m;
    }
    // both accumulators (general kernels: PathAcc; fast kernels: WalkAcc) must describe the same path
    PathAcc acc2; walk_acc_finish(wa, ei, ej, 0, 0, matches, acc2);
    acc2.A_last_match = acc.A_last_match; acc2.R_last_match = acc.R_last_match;     // (the fast kernels do not track these)
    if (memcmp(&acc, &acc2, sizeof acc) != 0 || pos2 != 0x00010001u) return PCB_ERR_RANGE;
    PathStats st; path_finalize(acc, head_h, head_v, tail_h, tail_v, m, n, st);
    out->read_start = st.read_start; out->read_end = st.read_end; out->adapter_start = st.adapter_start;
    out->adapter_end = st.adapter_end; out->raw_score = raw_score; out->matches = st.matches;
    out->aligned_len = st.aligned_len; out->full_len = st.full_len;
    return PCB_OK;
}

static std::mutex g_mu;
static pcb_context *g_ctx = nullptr;

// ------------------------------------------------------------------------------------------------
// check_compatibility (porechop_abi/ab_initio_src/compatibility.cpp:160-168): the second call site of the same DP in the
// reference -- candidate adapters against each other, linear gaps, match 2 / mismatch -1 / gap -1 (compatibility.h:9-12).
// ------------------------------------------------------------------------------------------------
extern "C" int pcb_check_compatibility(pcb_context *c, const char *blob, const int64_t *offsets, int32_t n_seqs,
                                       const int32_t *pair_a, const int32_t *pair_b, int64_t n_pairs, int32_t *flags)
{
    if (!c || n_seqs < 0 || n_pairs < 0 || (n_seqs > 0 && (!blob || !offsets)) || (n_pairs > 0 && (!pair_a || !pair_b || !flags))) return PCB_ERR_ARG;
    if (n_pairs == 0) return PCB_OK;
    if (c->deferred) return c->fail(PCB_ERR_ARG, "pcb_check_compatibility is not a deferred call");
    // String<Dna> (compatibility.h:24): a/c/g/t/u in either case, everything else becomes A
    const int64_t base = offsets[0], total = offsets[n_seqs] - base;
    std::string dna((size_t)total, 'A');
    for (int64_t k = 0; k < total; ++k) {
        const char ch = blob[base + k];
        if (strchr("ACGTUacgtu", ch) && ch) dna[(size_t)k] = ch;
    }
    std::vector<int64_t> off((size_t)n_seqs + 1), seg_off((size_t)n_pairs);
    for (int i = 0; i <= n_seqs; ++i) off[(size_t)i] = offsets[i] - base;
    std::vector<int32_t> seg_len((size_t)n_pairs), seg_ad((size_t)n_pairs), len1((size_t)n_pairs);
    for (int64_t p = 0; p < n_pairs; ++p) {
        int32_t a = pair_a[p], b = pair_b[p];
        if (a < 0 || a >= n_seqs || b < 0 || b >= n_seqs) return c->fail(PCB_ERR_ARG, "pair %lld: sequence index out of range", (long long)p);
        const int64_t la = off[(size_t)a + 1] - off[(size_t)a], lb = off[(size_t)b + 1] - off[(size_t)b];
        if (la == 0 || lb == 0) return c->fail(PCB_ERR_ARG, "pair %lld: empty sequence", (long long)p);
        if (la < lb) std::swap(a, b);                          // the longer sequence is row 0 (make_stringSet, compatibility.cpp:19-31)
        seg_off[(size_t)p] = off[(size_t)a]; seg_len[(size_t)p] = (int32_t)(off[(size_t)a + 1] - off[(size_t)a]); seg_ad[(size_t)p] = b;
        len1[(size_t)p] = seg_len[(size_t)p];
    }
    const PcbScores saved = c->sc;
    int rc = pcb_set_scoring(c, 2, -1, -1, -1);                // compatibility.h:9-12
    if (!rc) rc = pcb_set_adapters(c, dna.data(), off.data(), n_seqs);
    std::vector<pcb_result> rec((size_t)n_pairs);
    std::vector<uint8_t> aux((size_t)n_pairs);
    pcb_context::Slot &S = c->S();
    if (!rc && S.d_aux.reserve((size_t)n_pairs) != cudaSuccess) rc = c->fail(PCB_ERR_NOMEM, "cudaMalloc failed");
    S.want_aux = true;
    if (!rc) rc = pcb_align_tasks(c, dna.data(), total, seg_off.data(), seg_len.data(), seg_ad.data(), n_pairs, rec.data());
    S.want_aux = false;
    if (!rc && cudaMemcpy(aux.data(), S.d_aux.p, (size_t)n_pairs, cudaMemcpyDeviceToHost) != cudaSuccess) rc = c->fail(PCB_ERR_CUDA, "cudaMemcpy failed");
    c->sc = saved; c->stamp++;
    if (rc) return rc;
    for (int64_t p = 0; p < n_pairs; ++p) {
        const pcb_result &r = rec[(size_t)p];
        // get_flag (compatibility.cpp:121-156): errors are counted on [start, end) -- the last column is left out --, the
        // identity is an integer division
        const int mapped = r.aligned_len;
        const int dist = (r.aligned_len - 1) - (r.matches - (int)aux[(size_t)p]);
        const float identity = (float)((mapped - dist) * 100 / mapped);
        int flag = identity >= 87.5f ? 1 : 0;                  // compatibility.h:15
        const bool included = r.read_start > 3 || (len1[(size_t)p] - r.read_end) > 3;      // is_included, OVERLAP_LIMIT 3
        if (included && flag == 1) flag = 2;
        flags[p] = flag;
    }
    return PCB_OK;
}

extern "C" int check_compatibility(char *raw_seq1, char *raw_seq2)
{
    std::lock_guard<std::mutex> lock(g_mu);
    g_global_err.clear();
    int flag = 0;
    const auto failed = [&](const char *why) {
        g_global_err = why;
        fprintf(stderr, "cpp_functions.so (B200): check_compatibility failed: %s\n", why);
        if (g_dropin_fatal) exit(1);
        return 0;
    };
    if (!raw_seq1 || !raw_seq2) return failed("null sequence");
    if (!g_ctx) {
        int dev = 0;
        if (const char *e = getenv("PCB_DEVICE")) dev = atoi(e);
        if (pcb_create(dev, &g_ctx) != PCB_OK) return failed(g_global_err.c_str());
    }
    const std::string both = std::string(raw_seq1) + raw_seq2;
    const int64_t off[3] = { 0, (int64_t)strlen(raw_seq1), (int64_t)both.size() };
    const int32_t a = 0, b = 1;
    if (pcb_check_compatibility(g_ctx, both.data(), off, 2, &a, &b, 1, &flag) != PCB_OK) return failed(g_ctx->err.c_str());
    return flag;
}


static char *dup_cstring(const char *s)
{
    const size_t n = strlen(s);
    char *p = (char *)malloc(n + 1);      // released by freeCString, as adapter_align.cpp:34-44 does
    if (p) memcpy(p, s, n + 1);
    return p;
}

extern "C" char *adapterAlignment(char *readSeq, char *adapterSeq, int matchScore, int mismatchScore,
                                  int gapOpenScore, int gapExtensionScore)
{
    static const char *kFailed = "-1,0,-1,0,-2147483648,0.000000,0.000000";
    std::lock_guard<std::mutex> lock(g_mu);
    g_global_err.clear();
    if (!readSeq || !adapterSeq) { g_global_err = "null sequence"; return dup_cstring(kFailed); }
    if (!g_ctx) {
        int dev = 0;
        if (const char *e = getenv("PCB_DEVICE")) dev = atoi(e);
        if (pcb_create(dev, &g_ctx) != PCB_OK) {
            fprintf(stderr, "cpp_functions.so (B200): %s\n", g_global_err.c_str());
            if (g_dropin_fatal) exit(1);
            return dup_cstring(kFailed);
        }
    }
    pcb_context *c = g_ctx;
    const int64_t m = (int64_t)strlen(readSeq), n = (int64_t)strlen(adapterSeq);
    const int64_t aoff[2] = { 0, n };
    const int64_t soff = 0; const int32_t slen = (int32_t)m, aid = 0;
    pcb_result r;
    int rc = pcb_set_scoring(c, matchScore, mismatchScore, gapOpenScore, gapExtensionScore);
    if (!rc) rc = pcb_set_adapters(c, adapterSeq, aoff, 1);
    if (!rc) rc = pcb_align_tasks(c, readSeq, m, &soff, &slen, &aid, 1, &r);
    if (rc) {
        g_global_err = c->err;
        fprintf(stderr, "cpp_functions.so (B200): adapterAlignment failed: %s\n", c->err.c_str());
        if (g_dropin_fatal) exit(1);
        return dup_cstring(kFailed);
    }
    char buf[192];
    pcb_format_result(&r, buf, sizeof buf);
    return dup_cstring(buf);
}

extern "C" void freeCString(char *p) { free(p); }

// ------------------------------------------------------------------------------------------------
// pcb_align_windows_static: windows x ONE adapter without any allocation (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA graph capture, tight serving
// loops, harnesses that forbid it).  The two-pass trace of the fast classes (pcf_launch_pass1 / pcf_launch_pass2) runs with
// every byte of its working memory -- encoded windows, adapter codes, checkpoints, end cells, work items, sort tables,
// error flag -- carved out of ONE file-scope device array.  Every GPU that loads the library holds the array until the
// process exits; every call shares it and finds the last call's contents in it: what must start at zero is cleared by
// the call itself, on the caller's stream.  Two calls that overlap on one GPU would race, so a process-wide lock is held
// until the call's stream work is complete.
// ------------------------------------------------------------------------------------------------
#define PCB_STATIC_BYTES       ((unsigned)PCB_STATIC_SCRATCH_MB << 20)       /* limits: include/porechop_b200.h */
static __device__ __align__(256) unsigned char g_static_scratch[PCB_STATIC_BYTES];

namespace {
// byte offsets into g_static_scratch (each a multiple of 256)
constexpr size_t SO_ERR = 0, SO_AD_ASCII = 256, SO_AD_CODES = 512, SO_AD_OFF = 768, SO_IDS = 1024, SO_META = 1280,
                 SO_WINLEN = 4096, SO_ARENA = SO_WINLEN + sizeof(int32_t) * PCB_STATIC_MAX_WINDOWS,
                 SO_TRACE = SO_ARENA + (size_t)PCB_STATIC_MAX_WINDOWS * PCB_STATIC_MAX_STRIDE + 256;
static_assert(SO_TRACE % 256 == 0 && SO_TRACE < PCB_STATIC_BYTES, "layout of the static scratch");
static_assert((4 * PCF_NB + 1 + 16) * sizeof(int32_t) <= SO_WINLEN - SO_META, "sort tables of one adapter (pcf_meta_words(1))");

struct StaticAdapter { char seq[128]; int32_t n; };

// Device code may name the array directly.  Writes the adapter (ASCII, encoded by the next kernel), its offsets and id,
// and copies the window lengths, clamped to [0, stride] (error bit 128 when one was outside: the
// caller's array is device memory the host cannot check).
__global__ void k_static_setup(const StaticAdapter ad, const int32_t *win_len, int64_t n_windows, int32_t stride)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {
        int64_t *off = reinterpret_cast<int64_t *>(g_static_scratch + SO_AD_OFF);
        off[0] = 0; off[1] = ad.n;
        *reinterpret_cast<int32_t *>(g_static_scratch + SO_IDS) = 0;
    }
    if (i < 256) {
        g_static_scratch[SO_AD_ASCII + i] = i < ad.n ? (unsigned char)ad.seq[i] : (unsigned char)'A';
    }
    if (i < n_windows) {
        const int32_t l = win_len[i], lc = min(max(l, 0), stride);
        reinterpret_cast<int32_t *>(g_static_scratch + SO_WINLEN)[i] = lc;
        if (l != lc) atomicOr(reinterpret_cast<int32_t *>(g_static_scratch + SO_ERR), 128);
    }
}

struct StaticDev { unsigned char *ptr = nullptr; int grid1[PCF_NCLS] = {}, grid2[PCF_NCLS] = {}; };
StaticDev g_static_dev[64];
std::mutex g_static_mu;

int static_fail(int code, const char *fmt, ...)
{
    char buf[400];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_global_err = buf;
    return code;
}
#define PCB_CUDA_S(call) \
    do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { cudaGetLastError(); return static_fail(PCB_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e__)); } } while (0)
}  // namespace

extern "C" int pcb_align_windows_static(int device, const char *adapter, int32_t adapter_len, int match, int mismatch,
                                        int gap_open, int gap_extend, const char *d_windows, const int32_t *d_win_len,
                                        int64_t n_windows, int32_t stride, pcb_result *d_out, void *cuda_stream)
{
    if (!adapter || n_windows < 0 || (n_windows > 0 && (!d_windows || !d_win_len || !d_out)))
        return static_fail(PCB_ERR_ARG, "pcb_align_windows_static: null argument");
    if (n_windows > PCB_STATIC_MAX_WINDOWS || stride < 1 || stride > PCB_STATIC_MAX_STRIDE)
        return static_fail(PCB_ERR_RANGE, "pcb_align_windows_static takes at most %d windows of stride 1..%d per call: "
                           "got %lld windows of stride %d; its scratch is a fixed %u MB array", PCB_STATIC_MAX_WINDOWS,
                           PCB_STATIC_MAX_STRIDE, (long long)n_windows, stride, PCB_STATIC_BYTES >> 20);
    if (adapter_len < 1 || adapter_len > PCF_MAX_N2)
        return static_fail(PCB_ERR_RANGE, "pcb_align_windows_static takes adapters of 1..%d bases, got %d", PCF_MAX_N2, adapter_len);
    for (int i = 0; i < adapter_len; ++i)
        if (adapter[i] == 0 || !strchr("ACGTUacgtu", adapter[i]))
            return static_fail(PCB_ERR_RANGE, "pcb_align_windows_static takes adapters of A, C, G, T only (base %d is '%c')", i, adapter[i]);
    PcbScores sc; sc.match = match; sc.mismatch = mismatch; sc.gap_open = gap_open; sc.gap_extend = gap_extend; sc.linear = gap_open == gap_extend;
    const int cls = pcf_class_of_len(adapter_len);
    if (!fast_scheme(sc) || (pcf_class_two(cls) && !fast_scheme2(sc)))
        return static_fail(PCB_ERR_RANGE, "pcb_align_windows_static takes affine schemes with -7 <= gap_open < gap_extend < 0 < match only, got %d,%d,%d,%d",
                           match, mismatch, gap_open, gap_extend);
    if (device < 0 || device >= 64) return static_fail(PCB_ERR_ARG, "pcb_align_windows_static: device %d", device);
    if (n_windows == 0) return PCB_OK;

    std::lock_guard<std::mutex> lock(g_static_mu);             // calls share the array: one at a time, until its stream work is done
    PCB_CUDA_S(cudaSetDevice(device));
    StaticDev &D = g_static_dev[device];
    if (!D.ptr) {
        // set-up, once per device, with that device current: in host code the array's name is a stand-in, not an address
        unsigned char *p = nullptr;
        PCB_CUDA_S(cudaGetSymbolAddress((void **)&p, g_static_scratch));
        if (!p || ((uintptr_t)p & 255)) return static_fail(PCB_ERR_CUDA, "static scratch at %p", (void *)p);
        D.ptr = p;
    }
    if (!D.grid1[cls]) {                                        // occupancy query (also sets the kernels' shared-memory attribute on this device)
        const int g1 = pcf_occupancy_grid(1, cls), g2 = pcf_occupancy_grid(2, cls);
        if (g1 < 1 || g2 < 1) return static_fail(PCB_ERR_CUDA, "occupancy query failed for kernel class %d", cls);
        D.grid1[cls] = g1; D.grid2[cls] = g2;
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    unsigned char *const base = D.ptr;
    const uint64_t bytes = (uint64_t)n_windows * stride;

    PCB_CUDA_S(cudaMemsetAsync(base + SO_ERR, 0, 256, st));
    StaticAdapter ad; memset(&ad, 0, sizeof ad); memcpy(ad.seq, adapter, (size_t)adapter_len); ad.n = adapter_len;
    const unsigned gs = (unsigned)std::max<int64_t>(1, (n_windows + 255) / 256);
    k_static_setup<<<gs, 256, 0, st>>>(ad, d_win_len, n_windows, stride);
    PCB_CUDA_S(cudaGetLastError());
    PCB_CUDA_S(pcb_launch_encode(reinterpret_cast<const char *>(base + SO_AD_ASCII), base + SO_AD_CODES, 256, st));
    PCB_CUDA_S(pcb_launch_encode(d_windows, base + SO_ARENA, bytes, st));
    PCB_CUDA_S(cudaMemsetAsync(base + SO_ARENA + bytes, 0, 64, st));   // the kernels fetch bases a few bytes past a window's end

    PcbLaunch L; memset(&L, 0, sizeof L);
    L.ad.codes = base + SO_AD_CODES; L.ad.off = reinterpret_cast<const int64_t *>(base + SO_AD_OFF); L.ad.n = 1; L.ad.max_len = adapter_len;
    L.sc = sc;
    L.arena = base + SO_ARENA; L.arena_len = bytes;
    L.out = d_out; L.n_out = n_windows; L.err_flag = reinterpret_cast<int32_t *>(base + SO_ERR);
    L.src.n_windows = n_windows; L.src.stride = stride; L.src.win_len = reinterpret_cast<const int32_t *>(base + SO_WINLEN);
    L.src.ids = reinterpret_cast<const int32_t *>(base + SO_IDS); L.src.n_ids = 1; L.src.id_pos = nullptr; L.src.n_ids_total = 1;
    const int64_t n_units = (n_windows + 63) / 64;
    L.src.n_units = n_units;
    L.mmax = stride; L.smax = 1;
    L.meta = reinterpret_cast<int32_t *>(base + SO_META); L.n_major = 1;
    // trace scratch of a batch of units, as launch_dp lays it out: checkpoints | end cells | sorted items | surviving items
    const uint64_t per_unit = pcf_unit_words_trace(L.mmax, cls), ck_unit = per_unit - 128 - 1024;
    const uint64_t budget = (PCB_STATIC_BYTES - SO_TRACE) / sizeof(uint32_t);
    const int64_t per_batch = (int64_t)std::min<uint64_t>((uint64_t)n_units, budget / per_unit);
    if (per_batch < 1) return static_fail(PCB_ERR_RANGE, "pcb_align_windows_static: one unit of class %d needs %llu bytes of scratch", cls, (unsigned long long)per_unit * 4);
    uint32_t *const sp = reinterpret_cast<uint32_t *>(base + SO_TRACE);
    L.scratch = sp;
    L.entries = reinterpret_cast<uint4 *>(sp + ck_unit * (uint64_t)per_batch);
    L.sorted = reinterpret_cast<PcfItem *>(sp + (ck_unit + 128) * (uint64_t)per_batch);
    L.surv = reinterpret_cast<PcfItem *>(sp + (ck_unit + 128 + 512) * (uint64_t)per_batch);
    const int W = PCB_BLOCK / 32;
    for (int64_t u0 = 0; u0 < n_units; u0 += per_batch) {
        L.unit0 = u0; L.unit1 = std::min(n_units, u0 + per_batch);
        const int64_t ctas = (L.unit1 - L.unit0 + W - 1) / W;
        L.grid = (int)std::min<int64_t>(D.grid1[cls], ctas);
        PCB_CUDA_S(pcf_launch_pass1(L, cls, st));
        PCB_CUDA_S(pcf_launch_pass2(L, cls, (int)std::min<int64_t>(D.grid2[cls], ctas), st));
    }
    int32_t f = 0;
    PCB_CUDA_S(cudaMemcpyAsync(&f, base + SO_ERR, sizeof f, cudaMemcpyDeviceToHost, st));
    PCB_CUDA_S(cudaStreamSynchronize(st));
    if (f & 128) return static_fail(PCB_ERR_ARG, "pcb_align_windows_static: a window length outside [0, stride]");
    if (f) return static_fail(PCB_ERR_RANGE, "device reported alignment-window violation (flag %d)", f);
    return PCB_OK;
}
