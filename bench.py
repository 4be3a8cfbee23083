#!/usr/bin/env python3
"""
bench.py -- the headline benchmark of BASELINE.json: end + middle adapter search over synthetic ONT-like reads.

One "step" = one pass of the hot path over one batch of reads per GPU.  Workloads (--workload):
  ont        (default; BASELINE configs[2] cut to one GPU's share)
             (a) start windows (first 150 bases of every read) x the 119 start sequences,
                 end windows  (last 150 bases)             x the 117 end sequences      (phases A/B of the reference,
                 porechop_abi.py:200-245, :359-438; the adapter set has the shape of porechop_abi/adapters.py)
             (b) every read, trimmed by what (a) found, x 4 adapter sequences (phase C, porechop_abi.py:457-522)
  ultralong  (configs[3]) reads of 50-200 kb, full-length middle-adapter search only, length-balanced shards
  barcodes   (configs[4]) end windows x all 108 barcode sequences + kit adapters + 12 native "full sequence" sets
             (adapters.py:466-476), --require_two_barcodes binning decision per read inside the timed region
Every (window|read, adapter) pair is one adapterAlignment of the reference (src/adapter_align.cpp:11-31).

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run)
  python bench.py --impl reference ...                     (the reference's own CPU library, all host cores)

value   : reads/s, whole job over all ranks, inputs already resident in HBM (CUDA events, max over ranks)
e2e     : reads/s through the host-buffer C-ABI calls a caller makes (pinned host buffers; H2D of all inputs and D2H of
          all results inside the timed region; phase C is submitted only after phase B's results are on the host and
          the trims have been derived from them, as in the reference)
roofline: the dominant kernels (pass 1 + pass 2 of the trace of the 24-base class on the end windows), timed alone
"""
import argparse
import json
import os
import re
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

END_SIZE = 150
STRIDE = 160
N_MIDDLE_ADAPTERS = 4
EXTRA_END_TRIM = 2              # arg_parser.py:190 (--extra_end_trim)
END_THRESHOLD = 75.0            # arg_parser.py:192
MIN_TRIM_SIZE = 4               # arg_parser.py:187
BARCODE_THRESHOLD = 75.0        # arg_parser.py:156
BARCODE_DIFF = 5.0              # arg_parser.py:160
METRIC = 'end+middle adapter search throughput (reads/sec; GCUPS alongside)'
SCORES = (3, -6, -5, -2)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--workload', default='ont', choices=['ont', 'ultralong', 'barcodes'])
    ap.add_argument('--reads-per-gpu', type=int, default=0, help='0 = the workload\'s default')
    ap.add_argument('--seed', type=int, default=1)
    ap.add_argument('--cpu-sample-reads', type=int, default=0, help='reads in the CPU baseline sample (0 = auto)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-phase-e2e', action='store_true')
    args = ap.parse_args()
    if not args.reads_per_gpu:
        args.reads_per_gpu = {'ont': 131072, 'ultralong': 25000, 'barcodes': 131072}[args.workload]
    return args


# ----------------------------------------------------------------------------------------------------------------
# workloads
# ----------------------------------------------------------------------------------------------------------------
def barcode_adapter_set(seed=7):
    """Shape of a barcoded run's matching sets (porechop_abi.py:332-356): 96 forward + 12 reverse barcodes (24 bases,
    start + end sequence each), two kit adapter sets, and the 12 native 'full sequence' sets = flank + barcode + flank
    (adapters.py:466-476: 68-base start, 63-base end sequences)."""
    rng = np.random.default_rng(seed)
    letters = np.frombuffer(b'ACGT', dtype=np.uint8)
    rnd = lambda n: letters[rng.integers(0, 4, n)].tobytes()
    comp = bytes.maketrans(b'ACGT', b'TGCA')
    bc = [rnd(24) for _ in range(108)]
    bc_rc = [b.translate(comp)[::-1] for b in bc]
    kit_s, kit_e = [rnd(28), rnd(22)], [rnd(22), rnd(30)]
    f1, f2, f3, f4 = rnd(36), rnd(8), rnd(8), rnd(31)
    full_s = [f1 + bc[96 + i] + f2 for i in range(12)]
    full_e = [f3 + bc_rc[96 + i] + f4 for i in range(12)]
    seqs = bc + kit_s + full_s + bc_rc + kit_e + full_e
    n_s = len(bc) + len(kit_s) + len(full_s)
    start_ids = list(range(n_s)); end_ids = list(range(n_s, len(seqs)))
    is_barcode = np.zeros(n_s, dtype=np.uint8); is_barcode[:96] = 1      # the forward barcodes take part in the binning
    return seqs, start_ids, end_ids, is_barcode


def workload(args, rank, world):
    """This rank's share of the job: reads are dealt to the ranks by sharding.balanced_shards over the lengths of the
    whole job (the same list on every rank), then generated locally."""
    from custom_porechop_abi_b200 import synthetic
    from custom_porechop_abi_b200.sharding import balanced_shards
    n_total = args.reads_per_gpu * world
    rng = np.random.default_rng(args.seed)
    if args.workload == 'ultralong':
        lengths_all = rng.integers(50000, 200001, n_total).astype(np.int64)
    else:
        lengths_all = synthetic.read_lengths(n_total, rng)
    shards = balanced_shards(lengths_all, world)
    loads = np.array([lengths_all[s].sum() for s in shards], dtype=np.float64)
    mine = shards[rank]
    lengths = lengths_all[mine]
    n = len(mine)
    is_barcode = None
    if args.workload == 'barcodes':
        adapters, start_ids, end_ids, is_barcode = barcode_adapter_set()
    else:
        adapters, start_ids, end_ids = synthetic.adapter_set()
    blob, off = synthetic.make_reads(n, adapters, start_ids, end_ids, seed=args.seed + 1000 * rank, lengths=lengths)
    ad_len = np.array([len(a) for a in adapters], dtype=np.int64)
    wl = dict(kind=args.workload, adapters=adapters, ad_len=ad_len, start_ids=np.array(start_ids, dtype=np.int32),
              end_ids=np.array(end_ids, dtype=np.int32), blob=blob, off=off, n_reads=n, total_bases=int(off[-1]),
              lens=lengths.astype(np.int64), imbalance=float(loads.max() / loads.mean()), is_barcode=is_barcode,
              window_cells=0, task_cells=0)
    if args.workload != 'ultralong':
        sbuf, ebuf, wlen = synthetic.end_windows(blob, off, END_SIZE, STRIDE)
        wl.update(sbuf=sbuf, ebuf=ebuf, wlen=wlen)
        wl['window_cells'] = int(wlen.astype(np.int64).sum()) * int(ad_len[start_ids].sum() + ad_len[end_ids].sum())
    if args.workload != 'barcodes':
        mid_ids = np.array([start_ids[0], end_ids[0], start_ids[1], end_ids[1]], dtype=np.int32)[:N_MIDDLE_ADAPTERS]
        wl.update(mid_ids=mid_ids, seg_off=np.repeat(off[:-1], len(mid_ids)), seg_len=np.repeat(lengths.astype(np.int32), len(mid_ids)),
                  task_ad=np.tile(mid_ids, n))
        wl['task_cells'] = int(lengths.sum()) * int(ad_len[mid_ids].sum())
    wl['cells'] = wl['window_cells'] + wl['task_cells']
    return wl


def bench_config(args, wl=None):
    what = {'ont': 'synthetic ONT reads (log-normal lengths, N50 ~10 kb), %d reads per GPU: end-window alignment '
                   '(2 x %d-base windows vs 119 start + 117 end adapter sequences, shape of the Porechop database) '
                   '+ full-length middle-adapter search vs %d adapters' % (args.reads_per_gpu, END_SIZE, N_MIDDLE_ADAPTERS),
            'ultralong': 'synthetic ultra-long reads (uniform 50-200 kb), %d reads per GPU: full-length middle-adapter search vs '
                         '%d adapters, length-balanced shards' % (args.reads_per_gpu, N_MIDDLE_ADAPTERS),
            'barcodes': 'synthetic ONT reads, %d reads per GPU: end-window alignment (2 x %d-base windows) vs 108 barcode '
                        'sequences + kit adapters + 12 native full-sequence sets (63-68 bases) per end, --require_two_barcodes '
                        'binning per read' % (args.reads_per_gpu, END_SIZE)}[args.workload]
    cfg = {'workload': what, 'name': args.workload, 'reads_per_gpu': args.reads_per_gpu, 'end_size': END_SIZE,
           'scoring_scheme': list(SCORES),
           'l2': 'inputs + trace scratch per step exceed the 126 MB L2 (read arena and trace checkpoints are GB-sized)',
           'parallelism': 'reads dealt to the ranks by length-balanced shards (LPT), no data-path collective'}
    if args.workload == 'ont':
        cfg.update(adapter_sequences=236, middle_adapters=N_MIDDLE_ADAPTERS)
    return cfg


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (recipe in B200_PROFILING.md)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = False

    def run(self):
        q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
            'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'
        while not self.stop_flag:
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + q, '--format=csv,noheader,nounits'],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(',')])
            except Exception:
                pass
            time.sleep(0.15)

    def summary(self):
        sm = [int(s[0]) for s in self.samples if s[0].isdigit()]
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for k, n in enumerate(names) if any(len(s) > 2 + k and s[2 + k] == 'Active' for s in self.samples)]
        return {'sm_mhz': int(np.median(sm)) if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(sm)}


# ----------------------------------------------------------------------------------------------------------------
# CPU side: the reference's own library (oracle/_ref, compiled from /root/reference by oracle/Makefile) driven the way
# the reference drives it -- a thread pool of ctypes calls (porechop_abi.py:25,228-239) -- on a bounded sample.
# ----------------------------------------------------------------------------------------------------------------
def cpu_library():
    import ctypes
    from oracle_lib import REF_SO, oracle
    if os.path.isfile(REF_SO):
        lib = ctypes.CDLL(REF_SO)
        lib.adapterAlignment.argtypes = [ctypes.c_char_p, ctypes.c_char_p] + [ctypes.c_int] * 4
        lib.adapterAlignment.restype = ctypes.c_void_p
        lib.freeCString.argtypes = [ctypes.c_void_p]
        return 'reference', lib.adapterAlignment, lib.freeCString
    lib = oracle()
    return 'port', lib.pc_oracle_adapter_alignment, lib.pc_oracle_free


def cpu_step(wl, n_sample, threads, fn, free):
    """The step of the GPU arm restricted to the first n_sample reads. Returns (seconds, cells)."""
    from concurrent.futures import ThreadPoolExecutor
    adapters = wl['adapters']
    start_seqs = [adapters[i] for i in wl['start_ids']] if 'sbuf' in wl else []
    end_seqs = [adapters[i] for i in wl['end_ids']] if 'sbuf' in wl else []
    mid_seqs = [adapters[i] for i in wl['mid_ids']] if 'mid_ids' in wl else []
    blob, off = wl['blob'], wl['off']

    def one_read(r):
        if start_seqs:
            ws = bytes(wl['sbuf'][r, :wl['wlen'][r]]); we = bytes(wl['ebuf'][r, :wl['wlen'][r]])
            for a in start_seqs:
                free(fn(ws, a, *SCORES))
            for a in end_seqs:
                free(fn(we, a, *SCORES))
        if mid_seqs:
            full = bytes(blob[off[r]:off[r + 1]])
            for a in mid_seqs:
                free(fn(full, a, *SCORES))

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(one_read, range(n_sample), chunksize=4))
    dt = time.perf_counter() - t0
    ad_len = wl['ad_len']
    cells = 0
    if start_seqs:
        cells += int(wl['wlen'][:n_sample].astype(np.int64).sum()) * int(ad_len[wl['start_ids']].sum() + ad_len[wl['end_ids']].sum())
    if mid_seqs:
        cells += int(off[n_sample] - off[0]) * int(ad_len[wl['mid_ids']].sum())
    return dt, cells


def cpu_sample_size(wl, threads, target_s=15.0):
    # the reference's library does about 0.07 G cell updates / s / thread on this class of host
    per_read = wl['cells'] / wl['n_reads']
    return int(max(4, min(wl['n_reads'], target_s * 0.07e9 * threads / per_read)))


def run_reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    pool = {'ont': 4096, 'ultralong': 64, 'barcodes': 4096}[args.workload]
    small = argparse.Namespace(**vars(args)); small.reads_per_gpu = min(args.reads_per_gpu, pool)
    wl = workload(small, 0, 1)
    kind, fn, free = cpu_library()
    n_sample = args.cpu_sample_reads or cpu_sample_size(wl, threads, 8.0)
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_step(wl, max(4, n_sample // 8), threads, fn, free)
    times, cells = [], 0
    for _ in range(args.steps):
        dt, cells = cpu_step(wl, n_sample, threads, fn, free)
        times.append(dt)
    total = sum(times)
    value = n_sample * args.steps / total
    sample = '%d reads per step (same generator as the GPU arm), every alignment of the step through the library\'s adapterAlignment' % n_sample
    cfg = bench_config(args)
    cfg['sample_reads_per_step'] = n_sample
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'reads/s', 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1000.0 * total / args.steps, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'int32', 'data': 'synthetic', 'config': cfg,
            'gcups': cells * args.steps / total / 1e9,
            'cpu_baseline': {'value': value, 'unit': 'reads/s', 'cores': threads, 'kind': kind, 'sample': sample},
            'e2e': {'value': value, 'unit': 'reads/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line))


# ----------------------------------------------------------------------------------------------------------------
# host-side decisions of the e2e step (what the reference does between its phases)
# ----------------------------------------------------------------------------------------------------------------
def max_per_window(n, window, value):
    """out[w] = max(value over the hits of window w), 0 without hits; `window` ascending (as pcb_align_windows_hits returns)."""
    out = np.zeros(n, dtype=np.int64)
    if len(window):
        first = np.flatnonzero(np.r_[True, window[1:] != window[:-1]])
        out[window[first]] = np.maximum.reduceat(value.astype(np.int64), first)
    return out


def trims_from_hits(n, hs, he):
    """start_trim_amount / end_trim_amount of every read (nanopore_read.py:189, :211): max over the kept alignments."""
    a, b = hs.hits, he.hits
    start = max_per_window(n, a['window'], a['read_end'] + 1 + EXTRA_END_TRIM)
    end = max_per_window(n, b['window'], (END_SIZE - b['read_start']) + EXTRA_END_TRIM)
    return start, end


def trims_from_device(trim_s, trim_e):
    """The same from pcb_fetch_trims (the maximum is taken on the device: 4 bytes per read end, -1 = nothing kept)."""
    return np.where(trim_s >= 0, trim_s.astype(np.int64) + EXTRA_END_TRIM, 0), np.where(trim_e >= 0, trim_e.astype(np.int64) + EXTRA_END_TRIM, 0)


def two_barcode_calls(top_s, top_e):
    """determine_barcode with require_two_barcodes (nanopore_read.py:408-447) for every read at once, from the two best
    barcode identities per end (pcb_fetch_top2: descending, the first in adapter order among equals, as the reference's
    stable sort): both ends over the threshold, both ahead of their runner-up by barcode_diff, and the same barcode.
    -1 = 'none'."""
    def score(word):
        m = (word >> 9).astype(np.float64); length = (word & 0x1FF).astype(np.float64)
        with np.errstate(divide='ignore', invalid='ignore'):
            return np.where(length > 0, np.round(100.0 * m / length, 6), 0.0)      # the value parsed from "%f"
    s1, s2 = score(top_s['best_ident']), score(top_s['second_ident'])
    e1, e2 = score(top_e['best_ident']), score(top_e['second_ident'])
    ok = (s1 >= BARCODE_THRESHOLD) & (e1 >= BARCODE_THRESHOLD) & (s1 >= s2 + BARCODE_DIFF) & (e1 >= e2 + BARCODE_DIFF) & \
        (top_s['best'] == top_e['best']) & (top_s['best'] >= 0)
    return np.where(ok, top_s['best'], -1)


# ----------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the hot path has no CPU fallback')
    torch.cuda.set_device(local)
    numa_cpus = None
    if world > 1:                                # pinned buffers and the submitting thread next to this rank's GPU
        from custom_porechop_abi_b200.sharding import bind_to_gpu_cpus
        numa_cpus = bind_to_gpu_cpus(local)

    from custom_porechop_abi_b200.engine import AlignmentEngine, RESULT_DTYPE, HIT_DTYPE, TOP2_DTYPE
    from custom_porechop_abi_b200 import sharding

    wl = workload(args, rank, world)
    n = wl['n_reads']
    has_windows, has_tasks = 'sbuf' in wl, 'task_ad' in wl
    eng = AlignmentEngine(local)
    eng.set_adapters(wl['adapters'])

    keep = []

    def pin(a):
        a = np.ascontiguousarray(a)
        if a.dtype.names:                        # torch has no structured dtypes: pin as int32 words and view back
            t = torch.empty(a.shape + (a.dtype.itemsize // 4,), dtype=torch.int32).pin_memory()
            keep.append(t)
            v = t.numpy().view(a.dtype).reshape(a.shape)
            v[...] = a                               # (an uninitialised top2 buffer once made the checksum differ from run to run)
            return v
        if a.dtype == np.uint16:
            t = torch.empty(a.shape, dtype=torch.int16).pin_memory()
            keep.append(t)
            v = t.numpy().view(np.uint16); v[...] = a
            return v
        t = torch.from_numpy(a).pin_memory()
        keep.append(t)
        return t.numpy()

    if has_windows:
        sbuf, ebuf, wlen = pin(wl['sbuf']), pin(wl['ebuf']), pin(wl['wlen'])
        want_s = wl['is_barcode'] if wl['is_barcode'] is not None else (wl['ad_len'][wl['start_ids']] == 24).astype(np.uint8)
        want_e = wl['is_barcode'] if wl['is_barcode'] is not None else (wl['ad_len'][wl['end_ids']] == 24).astype(np.uint8)
        hits_s, hits_e = pin(np.empty(16 * n, dtype=HIT_DTYPE)), pin(np.empty(16 * n, dtype=HIT_DTYPE))
        ident_s = pin(np.zeros((n, int(want_s.sum())), dtype=np.uint16))
        ident_e = pin(np.zeros((n, int(want_e.sum())), dtype=np.uint16))
        top_s, top_e = pin(np.zeros(n, dtype=TOP2_DTYPE)), pin(np.zeros(n, dtype=TOP2_DTYPE))
        trim_s, trim_e = pin(np.full(n, -1, dtype=np.int32)), pin(np.full(n, -1, dtype=np.int32))
    if has_tasks:
        blob, task_ad = pin(wl['blob']), pin(wl['task_ad'])
        seg_off, seg_len = pin(wl['seg_off']), pin(wl['seg_len'])
        out_m = pin(np.empty(len(task_ad), dtype=RESULT_DTYPE))
    lens64, off64 = wl['lens'], wl['off'][:-1]

    def barrier():
        eng.synchronize()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device='cuda')
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- resident arm: inputs staged once, timed region = the kernels (full 32-byte records stay in HBM)
    if has_windows:
        eng.select_slot(0); eng.stage_windows(sbuf, wlen, wl['start_ids'])
        eng.select_slot(1); eng.stage_windows(ebuf, wlen, wl['end_ids'])
    if has_tasks:
        eng.select_slot(2); eng.stage_tasks(blob, seg_off, seg_len, task_ad)

    def resident_step():
        if has_windows:
            eng.select_slot(0); eng.run_windows_resident()
            eng.select_slot(1); eng.run_windows_resident()
        if has_tasks:
            eng.select_slot(2); eng.run_tasks_resident()

    for _ in range(args.warmup):
        resident_step()
    barrier()
    eng.reset_counters()
    sampler = ClockSampler(local); sampler.start()
    eng.timer_start()
    for _ in range(args.steps):
        resident_step()
    ms_total = eng.timer_stop()
    barrier()
    cnt = eng.counters()
    ms_total = max_over_ranks(ms_total)
    ms_step = ms_total / args.steps

    # ---- end-to-end arm: the host-buffer C-ABI calls a user makes; H2D of the inputs and D2H of the results every step.
    #      Phase B comes back reduced (pcb_align_windows_hits); phase C is submitted only after phase B's results are on
    #      the host and the trims have been derived from them (nanopore_read.py:66-72, :225)
    state = {}
    binning = wl['kind'] == 'barcodes'

    # the batch goes through in PARTS parts (the reference's pool works through the reads in order, too): phase C of a part
    # is queued as soon as that part's phase B is on the host, beside phase B of the next part
    PARTS = 2 if (has_windows and has_tasks) else 1
    cut = [n * p // PARTS for p in range(PARTS + 1)]
    A = N_MIDDLE_ADAPTERS

    stamps = []

    def e2e_step(trace=False):
        mark = (lambda what: stamps.append((what, time.perf_counter()))) if trace else (lambda what: None)
        mark('start')
        if has_tasks and has_windows:
            # the reads are all in host memory before the first alignment (porechop_abi.py:61-71): announce the blob, so
            # that it crosses PCIe under the window kernels; every byte still travels inside the timed region, every step
            eng.upload_reads(blob)
        res = []
        if has_windows:
            for p in range(PARTS):
                a, b = cut[p], cut[p + 1]
                eng.select_slot(2 * p)
                hs = eng.align_windows_hits(sbuf[a:b], wlen[a:b], wl['start_ids'], END_THRESHOLD, False, END_SIZE, MIN_TRIM_SIZE, want_ident=want_s,
                                            hits_buf=hits_s[16 * a:16 * b], ident_buf=False if binning else ident_s[a:b])
                if binning:
                    eng.fetch_top2(b - a, out=top_s[a:b])
                if has_tasks:
                    eng.fetch_trims(b - a, out=trim_s[a:b])
                eng.select_slot(2 * p + 1)
                he = eng.align_windows_hits(ebuf[a:b], wlen[a:b], wl['end_ids'], END_THRESHOLD, True, END_SIZE, MIN_TRIM_SIZE, want_ident=want_e,
                                            hits_buf=hits_e[16 * a:16 * b], ident_buf=False if binning else ident_e[a:b])
                if binning:
                    eng.fetch_top2(b - a, out=top_e[a:b])
                if has_tasks:
                    eng.fetch_trims(b - a, out=trim_e[a:b])
                res.append((hs, he))
                mark('B%d queued' % p)
        n_hits = [0, 0]
        for p in range(PARTS):
            a, b = cut[p], cut[p + 1]
            if has_windows:
                hs, he = res[p]
                if PARTS > 1:
                    eng.wait_slot(2 * p); eng.wait_slot(2 * p + 1)      # this part's phase B is on the host; the rest keeps running
                else:
                    eng.wait()
                n_hits[0] += hs.n_hits; n_hits[1] += he.n_hits
                mark('B%d on the host' % p)
            if has_tasks:
                if has_windows:          # the trimmed reads are what phase C aligns (nanopore_read.py:66-72, :225)
                    st, et = trims_from_device(trim_s[a:b], trim_e[a:b])
                    lo = np.minimum(st, lens64[a:b])
                    hi = np.maximum(lens64[a:b] - et, lo)
                    seg_off[A * a:A * b] = np.repeat(off64[a:b] + lo, A)
                    seg_len[A * a:A * b] = np.repeat(hi - lo, A)
                eng.select_slot(4 + p)
                mark('trims %d' % p)
                # (without window calls to hide it behind, the blob goes with the task call itself: in pieces, each with its own
                #  plan, so that the kernels of one piece run under the copy of the next)
                eng.align_tasks(None if has_windows else blob, seg_off[A * a:A * b], seg_len[A * a:A * b], task_ad[A * a:A * b], out=out_m[A * a:A * b])
                mark('C%d queued' % p)
                if trace and has_windows:                   # (the untimed trace step also checks the device's trims against the hit lists)
                    st2, et2 = trims_from_hits(b - a, hs, he)
                    assert np.array_equal(st, st2) and np.array_equal(et, et2)
        eng.wait()
        mark('all collected')
        state['hits'] = tuple(n_hits)
        state['parts'] = res
        if binning:
            state['calls'] = two_barcode_calls(top_s, top_e)

    eng.set_deferred(True)
    for _ in range(max(1, min(args.warmup, 2))):
        e2e_step()
    barrier()
    c0 = eng.counters()
    t0 = time.perf_counter()
    eng.timer_start()
    for _ in range(args.steps):
        e2e_step()
    e2e_ms_ev = eng.timer_stop()
    e2e_s = max(time.perf_counter() - t0, e2e_ms_ev / 1000.0)
    barrier()
    c1 = eng.counters()
    if os.environ.get('PCB_E2E_TRACE') and rank == 0:          # host-side time line of one more (untimed) step, to stderr
        e2e_step(trace=True)
        print('e2e step: ' + ', '.join('%s %.1f ms' % (w_, (t_ - stamps[0][1]) * 1e3) for w_, t_ in stamps[1:]), file=sys.stderr)
        for rep in range(4):                                   # ... and the parts of the checksum, step after step (they must not move)
            e2e_step()
            parts_sum = {}
            if has_tasks:
                for f_ in ('raw_score', 'read_start', 'read_end', 'matches'):
                    parts_sum['C ' + f_] = int(out_m[f_].astype(np.int64).sum())
            if has_windows:
                parts_sum['hits'] = [int(h.hits['raw_score'].astype(np.int64).sum()) for pair in state['parts'] for h in pair]
                parts_sum['ident'] = [int(ident_s.astype(np.int64).sum()), int(ident_e.astype(np.int64).sum())]
                parts_sum['trim'] = [int(trim_s.astype(np.int64).sum()), int(trim_e.astype(np.int64).sum())]
            print('e2e parts %d: %s' % (rep, parts_sum), file=sys.stderr)
    eng.set_deferred(False)
    sampler.stop_flag = True; sampler.join(timeout=3)
    e2e_s = max_over_ranks(e2e_s)
    checksum = 0
    if has_tasks:
        checksum += int(out_m['raw_score'].astype(np.int64).sum())
    if has_windows:
        checksum += sum(int(h.hits['raw_score'].astype(np.int64).sum()) for pair in state['parts'] for h in pair)
        checksum += int(ident_s.astype(np.int64).sum() + ident_e.astype(np.int64).sum()) + int(top_s['best_ident'].astype(np.int64).sum())
    # the summary counters of the whole job, merged over NCCL as the reference sums them in-process (porechop_abi.py:441-454)
    local_counts = [n, state.get('hits', (0, 0))[0], state.get('hits', (0, 0))[1]]
    if 'calls' in state:
        local_counts.append(int((state['calls'] >= 0).sum()))
    merged_counts = sharding.merge_counters(local_counts, device='cuda').tolist()
    n_job = int(sum_over_ranks(n))
    cells_job = sum_over_ranks(wl['cells'])

    # ---- dominant kernels alone: the trace of the 24-base class on the end windows (pass 1 + pass 2), CUDA events on the
    #      library's own stream; and the scan of the full-length search with its planner / merge / window-trace launches
    ad_len = wl['ad_len']
    kern = None
    if has_windows:
        ids24 = np.array([i for i in wl['end_ids'] if ad_len[i] == 24], dtype=np.int32)
        eng.select_slot(3); eng.stage_windows(ebuf, wlen, ids24)
        eng.run_windows_resident(); eng.synchronize()
        reps = 5
        c_k0 = eng.counters()
        eng.timer_start()
        for _ in range(reps):
            eng.run_windows_resident()
        k_ms = eng.timer_stop() / reps
        kern = dict(ms=k_ms, launches=(eng.counters()['dp_kernel_launches'] - c_k0['dp_kernel_launches']) // reps,
                    cells=int(wlen.astype(np.int64).sum()) * int(ad_len[ids24].sum()), tasks=n * len(ids24), n_ids=len(ids24),
                    # algorithmic bytes of that launch: window bases read once + adapters + one 32-byte record per pair
                    bytes=int(wlen.astype(np.int64).sum()) + int(ad_len[ids24].sum()) + 32 * n * len(ids24))
    scan = None
    if has_tasks:
        eng.select_slot(2)
        eng.stage_tasks(blob, np.repeat(off64, N_MIDDLE_ADAPTERS), np.repeat(lens64.astype(np.int32), N_MIDDLE_ADAPTERS), task_ad)
        eng.run_tasks_resident(); eng.synchronize()
        reps = 3
        eng.timer_start()
        for _ in range(reps):
            eng.run_tasks_resident()
        scan = dict(ms=eng.timer_stop() / reps, cells=wl['task_cells'], bytes=wl['total_bases'] + 32 * len(task_ad))

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))
        except Exception:
            pass
        hbm_peak = float(peaks.get('hbm_gbs', 6650.0))
        peak_src = 'measured (MEASURED_PEAKS.json)' if 'hbm_gbs' in peaks else 'fallback (B200_PROFILING.md)'
        clocks = sampler.summary()
        sm_mhz = clocks.get('sm_mhz') or 1965
        # integer-issue roofline: a DP cell pair (two tasks packed in one register) needs at least 4
        # integer warp instructions (diagonal add, H, V, max3); the pipe micro-benchmark (scripts/microbench/pipes.cu ->
        # profiles/pipes_*.txt) gives the clocks per integer warp instruction and sub-partition
        instr_per_pair = 4.0
        clk_mixed, clk_dpx, pipes_file = pipe_clocks()
        int_peak = 148 * 4 * 64 / (instr_per_pair * clk_mixed) * sm_mhz * 1e6 / 1e9
        strict_peak = 148 * 4 * 64 / (3.0 * clk_dpx) * sm_mhz * 1e6 / 1e9
        traffic = profile_traffic()
        dom = kern if kern is not None else scan
        dom_name = ('k_fast<pass 1,24,EXACT> + k_rounds<24,EXACT> (+ sort kernels): end windows x %d 24-base end sequences, %d launches dealt to the library\'s 4 streams'
                    % (kern['n_ids'], kern['launches'])) if kern is not None else 'k_fast<scan,24> + planner / merge / window traces (full-length search)'
        line = {
            'metric': METRIC, 'value': n_job / (ms_step / 1000.0), 'unit': 'reads/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_step, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'int16x2', 'data': 'synthetic',
            'config': bench_config(args, wl),
            'shard_imbalance': wl['imbalance'],
            'host_binding': ('rank bound to the %d CPUs next to its GPU (NVML affinity)' % len(numa_cpus)) if numa_cpus else 'none',
            'gcups': cells_job / (ms_step / 1000.0) / 1e9,
            'e2e': {'value': n_job * args.steps / e2e_s, 'unit': 'reads/s',
                    'h2d_bytes_per_step': int((c1['h2d_bytes'] - c0['h2d_bytes']) / args.steps),
                    'd2h_bytes_per_step': int((c1['d2h_bytes'] - c0['d2h_bytes']) / args.steps),
                    'gcups': cells_job * args.steps / e2e_s / 1e9, 'checksum': checksum,
                    'phases': 'in %d parts: B (reduced: hits, trim amounts, 16-bit barcode identities) -> that part collected on the host -> C on the trimmed reads (32-byte records) beside the next part\'s B' % PARTS
                              if has_windows and has_tasks else ('B (reduced: hits + the two best barcode identities per read and end) -> wait -> binning decision on the host' if has_windows else 'C'),
                    'job_counters': {'reads': merged_counts[0], 'start_hits': merged_counts[1], 'end_hits': merged_counts[2],
                                     **({'reads_binned': merged_counts[3]} if len(merged_counts) > 3 else {})}},
            'gpu_launches': int(cnt['kernel_launches']),
            'clocks': clocks,
            'roofline': {'bound': 'int_issue', 'kernel': dom_name,
                         'achieved': dom['cells'] / (dom['ms'] / 1000.0) / 1e9, 'peak': int_peak, 'unit': 'GCUPS',
                         'frac': dom['cells'] / (dom['ms'] / 1000.0) / 1e9 / int_peak,
                         'peak_model': {'formula': '148 SM x 4 sub-partitions x 64 cells / (instr_per_pair x clk_per_instr) x SM clock',
                                        'instr_per_pair': instr_per_pair, 'clk_per_instr': clk_mixed, 'source': pipes_file,
                                        'strict_peak': strict_peak, 'strict_model': '3 DPX instructions per cell pair on the ALU pipe alone at %.3f clk' % clk_dpx,
                                        'strict_frac': dom['cells'] / (dom['ms'] / 1000.0) / 1e9 / strict_peak},
                         'launch_ms': dom['ms'], 'traffic': traffic.get('bytes_per_pair') and int(traffic['bytes_per_pair'] * dom.get('tasks', 0)) or None,
                         'traffic_source': traffic.get('source'),
                         'hbm': {'achieved': dom['bytes'] / (dom['ms'] / 1000.0) / 1e9, 'peak': hbm_peak, 'unit': 'GB/s',
                                 'frac': dom['bytes'] / (dom['ms'] / 1000.0) / 1e9 / hbm_peak, 'peak_source': peak_src,
                                 'note': 'algorithmic bytes only (bases once, one 32-byte record per pair): the kernels are bound by integer issue, not HBM'}},
            'int_dp': {'kernel_gcups': kern and kern['cells'] / (kern['ms'] / 1000.0) / 1e9, 'model_peak_gcups': int_peak,
                       'frac': kern and kern['cells'] / (kern['ms'] / 1000.0) / 1e9 / int_peak,
                       'scan_phase_gcups': scan and scan['cells'] / (scan['ms'] / 1000.0) / 1e9,
                       'scan_phase_frac': scan and scan['cells'] / (scan['ms'] / 1000.0) / 1e9 / int_peak, 'sm_mhz': sm_mhz},
        }
        if world == 1 and not args.no_phase_e2e and args.workload == 'ont':
            line['phase_e2e'] = phase_e2e(eng, wl, blob)
            if os.environ.get('PCB_PHASE_DEBUG'):              # a second time, and on an engine of its own
                phase_e2e(eng, wl, blob)
                eng2 = AlignmentEngine(local)
                phase_e2e(eng2, wl, blob)
                eng2.close()
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            kind, fn, free = cpu_library()
            ns = args.cpu_sample_reads or cpu_sample_size(wl, threads, 15.0)
            dt, cells = cpu_step(wl, ns, threads, fn, free)
            line['cpu_baseline'] = {'value': ns / dt, 'unit': 'reads/s', 'cores': threads, 'kind': kind, 'gcups': cells / dt / 1e9,
                                    'sample': 'first %d reads of the same batch, one pass, %d threads of ctypes calls' % (ns, threads)}
        print(json.dumps(line))
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def pipe_clocks():
    """clk per integer warp instruction and sub-partition from the newest profiles/pipes_*.txt: (ALU + FMA pipes alternating,
    DPX alone), at 4 warps per sub-partition."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', 'pipes_*.txt')))
    mixed, dpx, name = 1.8, 2.06, 'fallback constants (no profiles/pipes_*.txt)'
    if files:
        name = 'profiles/' + os.path.basename(files[-1])
        for line in open(files[-1]):
            m = re.match(r'(\S.*?)\s+warps/SMSP=4\s+clk/warp-instr/SMSP =\s*([0-9.]+)', line)
            if m and m.group(1).strip() == 'VIADDMNMX + IMAD(rr,imm)':
                mixed = float(m.group(2))
            if m and m.group(1).strip() == 'VIADDMNMX.S16x2':
                dpx = float(m.group(2))
    return mixed, dpx, name


def profile_traffic():
    """DRAM bytes per (window, adapter) pair of the dominant kernels from the newest profiles/traffic_*.json (written by
    scripts/ncu_summary.py from an ncu --set full capture); nothing when there is none."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', 'traffic_*.json')))
    if not files:
        return {}
    try:
        d = json.load(open(files[-1]))
        return {'bytes_per_pair': float(d['dram_bytes']) / float(d['pairs']), 'source': 'profiles/' + os.path.basename(files[-1]) + ': ' + d.get('what', '')}
    except Exception:
        return {}


def phase_e2e(eng, wl, pinned_blob, n_reads=0, reps=3):
    """reads/s through the reference-facing phase functions (porechop_batch.find_*) on reference-shaped read objects built
    from a ReadBlob: what a maintainer who swaps the three loops gets, Python included.  One untimed pass (buffers of the
    library grow to the job's size, as the bench's own warm-up steps do), then `reps` timed passes; every pass builds its
    own ReadBlob and read objects and sends the read bases to the device again."""
    from custom_porechop_abi_b200 import porechop_batch as pb
    from custom_porechop_abi_b200.read_blob import ReadBlob
    from custom_porechop_abi_b200 import synthetic
    n = min(n_reads or wl['n_reads'], wl['n_reads'])
    sets = synthetic.adapter_sets(wl['adapters'], wl['start_ids'], wl['end_ids'])
    passes = []
    import gc
    for rep in range(reps + 1):
        gc.collect()            # every pass starts like a fresh job (a full collection of the last pass's 131,072 objects once landed in phase A: 81 instead of 24 ms)
        h2d0 = eng.counters()['h2d_bytes']
        t_make = time.perf_counter()
        rb = ReadBlob.from_buffer(pinned_blob[:wl['off'][n]], wl['off'][:n + 1])  # the reads as loaded: one (pinned) buffer + offsets
        reads = rb.make_reads()
        eng._uploaded = None                                                      # nothing of an earlier pass counts as resident
        t0 = time.perf_counter()
        matching = pb.find_matching_adapter_sets(reads[:10000], 0, END_SIZE, list(SCORES), None, 90.0, 1, sets, engine=eng, read_blob=rb.head(10000))
        t1 = time.perf_counter()
        pb.find_adapters_at_read_ends(reads, matching or sets[:2], 0, END_SIZE, EXTRA_END_TRIM, END_THRESHOLD, list(SCORES), None, MIN_TRIM_SIZE,
                                      1, False, BARCODE_THRESHOLD, BARCODE_DIFF, False, 'forward', engine=eng, read_blob=rb)
        t2 = time.perf_counter()
        pb.find_adapters_in_read_middles(reads, matching or sets[:2], 0, 90.0, 10, 100, list(SCORES), None, 1, False, engine=eng, read_blob=rb)
        t3 = time.perf_counter()
        passes.append({'objects': t0 - t_make, 'phase_a': t1 - t0, 'phase_b': t2 - t1, 'phase_c': t3 - t2, 'total': t3 - t0,
                       'h2d_bytes': int(eng.counters()['h2d_bytes'] - h2d0)})
        del reads, rb
    if os.environ.get('PCB_PHASE_DEBUG'):
        for k_, p_ in enumerate(passes):
            print('phase_e2e pass %d: %s' % (k_, {a_: round(b_ * 1e3, 1) if a_ != 'h2d_bytes' else b_ for a_, b_ in p_.items()}), file=sys.stderr)
    timed = passes[1:]
    total = sum(p['total'] for p in timed)
    mean = {k: sum(p[k] for p in timed) / len(timed) for k in ('objects', 'phase_a', 'phase_b', 'phase_c')}
    return {'value': n * len(timed) / total, 'unit': 'reads/s', 'reads': n, 'passes': len(timed), 'warmup_passes': 1,
            'matching_sets': len(matching), 'seconds': mean, 'first_pass_seconds': passes[0]['total'],
            'h2d_bytes_per_pass': timed[-1]['h2d_bytes'],
            'what': 'porechop_batch.find_matching_adapter_sets (10,000 check reads x 236 sequences) + find_adapters_at_read_ends + '
                    'find_adapters_in_read_middles on NanoporeRead-shaped objects over a ReadBlob, matching sets as phase A finds them; '
                    'mean of the timed passes, read bases sent to the device in every pass'}


def main():
    # the contract is ONE JSON line on stdout: libraries that print there (NCCL's version banner ...) go to stderr instead
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, 'w')
    args = parse_args()
    if args.impl == 'reference':
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == '__main__':
    main()
