#!/usr/bin/env python3
"""
bench.py -- the headline benchmark of BASELINE.json: end + middle adapter search over synthetic ONT-like reads.

One "step" = one pass of the hot path over one batch of reads per GPU:
  (a) start windows (first 150 bases of every read) x the 119 start sequences,
      end windows  (last 150 bases)             x the 117 end sequences      (phases A/B of the reference,
      porechop_abi.py:200-245, :359-438; the adapter set has the shape of porechop_abi/adapters.py)
  (b) every full-length read x 4 adapter sequences (phase C, porechop_abi.py:457-522)
Every (window|read, adapter) pair is one adapterAlignment of the reference (src/adapter_align.cpp:11-31).

  python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run)
  python bench.py --impl reference ...                     (the reference's own CPU library, all host cores)

value   : reads/s, whole job over all ranks, inputs already resident in HBM (CUDA events, max over ranks)
e2e     : reads/s through the host-buffer C-ABI calls (pinned host buffers, H2D + D2H inside the timed region)
roofline: the dominant kernel (k_fast<TRACE,24,EXACT> on the end windows), timed alone with CUDA events
"""
import argparse
import json
import os
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
METRIC = 'end+middle adapter search throughput (reads/sec; GCUPS alongside)'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--reads-per-gpu', type=int, default=131072)
    ap.add_argument('--seed', type=int, default=1)
    ap.add_argument('--cpu-sample-reads', type=int, default=0, help='reads in the CPU baseline sample (0 = auto)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    return ap.parse_args()


def workload(n_reads, seed):
    from custom_porechop_abi_b200 import synthetic
    adapters, start_ids, end_ids = synthetic.adapter_set()
    blob, off = synthetic.make_reads(n_reads, adapters, start_ids, end_ids, seed=seed)
    sbuf, ebuf, wlen = synthetic.end_windows(blob, off, END_SIZE, STRIDE)
    mid_ids = np.array([start_ids[0], end_ids[0], start_ids[1], end_ids[1]], dtype=np.int32)[:N_MIDDLE_ADAPTERS]
    lens = (off[1:] - off[:-1]).astype(np.int32)
    seg_off = np.repeat(off[:-1], len(mid_ids))
    seg_len = np.repeat(lens, len(mid_ids))
    task_ad = np.tile(mid_ids, n_reads)
    ad_len = np.array([len(a) for a in adapters], dtype=np.int64)
    cells = int(wlen.astype(np.int64).sum()) * int(ad_len[start_ids].sum() + ad_len[end_ids].sum()) + \
        int(lens.astype(np.int64).sum()) * int(ad_len[mid_ids].sum())
    return dict(adapters=adapters, start_ids=np.array(start_ids, dtype=np.int32), end_ids=np.array(end_ids, dtype=np.int32),
                blob=blob, off=off, sbuf=sbuf, ebuf=ebuf, wlen=wlen, seg_off=seg_off, seg_len=seg_len, task_ad=task_ad,
                mid_ids=mid_ids, cells=cells, n_reads=n_reads, total_bases=int(off[-1]),
                window_cells=int(wlen.astype(np.int64).sum()) * int(ad_len[start_ids].sum() + ad_len[end_ids].sum()))


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
    from oracle_lib import REF_SO, ORACLE_SO, oracle
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
    sc = (3, -6, -5, -2)
    start_seqs = [adapters[i] for i in wl['start_ids']]
    end_seqs = [adapters[i] for i in wl['end_ids']]
    mid_seqs = [adapters[i] for i in wl['mid_ids']]
    blob, off, sbuf, ebuf, wlen = wl['blob'], wl['off'], wl['sbuf'], wl['ebuf'], wl['wlen']

    def one_read(r):
        ws = bytes(sbuf[r, :wlen[r]]); we = bytes(ebuf[r, :wlen[r]])
        full = bytes(blob[off[r]:off[r + 1]])
        for a in start_seqs:
            free(fn(ws, a, *sc))
        for a in end_seqs:
            free(fn(we, a, *sc))
        for a in mid_seqs:
            free(fn(full, a, *sc))

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(one_read, range(n_sample), chunksize=4))
    dt = time.perf_counter() - t0
    ad_len = np.array([len(a) for a in adapters], dtype=np.int64)
    cells = int(wlen[:n_sample].astype(np.int64).sum()) * int(ad_len[wl['start_ids']].sum() + ad_len[wl['end_ids']].sum()) + \
        int(off[n_sample] - off[0]) * int(ad_len[wl['mid_ids']].sum())
    return dt, cells


def cpu_sample_size(wl, threads, target_s=15.0):
    # the reference's library does about 0.07 G cell updates / s / thread on this class of host
    per_read = wl['cells'] / wl['n_reads']
    return int(max(8, min(wl['n_reads'], target_s * 0.07e9 * threads / per_read)))


def run_reference_arm(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    wl = workload(min(args.reads_per_gpu, 4096), args.seed)
    kind, fn, free = cpu_library()
    n_sample = args.cpu_sample_reads or cpu_sample_size(wl, threads, 8.0)
    for _ in range(max(0, min(args.warmup, 1))):
        cpu_step(wl, max(8, n_sample // 8), threads, fn, free)
    times, cells = [], 0
    for _ in range(args.steps):
        dt, cells = cpu_step(wl, n_sample, threads, fn, free)
        times.append(dt)
    total = sum(times)
    value = n_sample * args.steps / total
    sample = '%d reads per step (same generator and seed as the GPU arm): 2 windows x 236 adapters + full read x %d adapters' % (
        n_sample, N_MIDDLE_ADAPTERS)
    line = {'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'reads/s', 'n_gpus': args.gpus, 'steps': args.steps,
            'warmup': args.warmup, 'ms_per_step': 1000.0 * total / args.steps, 'higher_is_better': True, 'scaling': 'weak',
            'vs_baseline': None, 'dtype': 'int32', 'data': 'synthetic',
            'config': bench_config(args, wl),
            'gcups': cells * args.steps / total / 1e9,
            'cpu_baseline': {'value': value, 'unit': 'reads/s', 'cores': threads, 'kind': kind, 'sample': sample},
            'e2e': {'value': value, 'unit': 'reads/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line))


def bench_config(args, wl):
    return {'workload': 'synthetic ONT reads (log-normal lengths, N50 ~10 kb), %d reads per GPU: end-window alignment '
                        '(2 x %d-base windows vs 119 start + 117 end adapter sequences, shape of the Porechop database) '
                        '+ full-length middle-adapter search vs %d adapters' % (args.reads_per_gpu, END_SIZE, N_MIDDLE_ADAPTERS),
            'reads_per_gpu': args.reads_per_gpu, 'end_size': END_SIZE, 'adapter_sequences': 236,
            'middle_adapters': N_MIDDLE_ADAPTERS, 'scoring_scheme': [3, -6, -5, -2],
            'l2': 'inputs + trace scratch per step exceed the 126 MB L2 (read arena and trace ring are GB-sized)',
            'parallelism': 'reads sharded across ranks, no data-path collective'}


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

    from custom_porechop_abi_b200.engine import AlignmentEngine, RESULT_DTYPE, _ptr
    from custom_porechop_abi_b200 import cpp_function_wrappers as w

    wl = workload(args.reads_per_gpu, args.seed + 1000 * rank)
    n = wl['n_reads']
    eng = AlignmentEngine(local)
    eng.set_adapters(wl['adapters'])

    def pinned(a):
        a = np.ascontiguousarray(a)
        if a.dtype == RESULT_DTYPE:            # torch has no structured dtypes: pin as int32 [..., 8] and view back
            t = torch.empty(a.shape + (8,), dtype=torch.int32).pin_memory()
            return t, t.numpy().view(RESULT_DTYPE).reshape(a.shape)
        t = torch.from_numpy(a).pin_memory()
        return t, t.numpy()

    keep = []
    def pin(a):
        t, v = pinned(a); keep.append(t); return v
    sbuf, ebuf, wlen = pin(wl['sbuf']), pin(wl['ebuf']), pin(wl['wlen'])
    blob, seg_off, seg_len, task_ad = pin(wl['blob']), pin(wl['seg_off']), pin(wl['seg_len']), pin(wl['task_ad'])
    out_s = pin(np.empty((n, len(wl['start_ids'])), dtype=RESULT_DTYPE))
    out_e = pin(np.empty((n, len(wl['end_ids'])), dtype=RESULT_DTYPE))
    out_m = pin(np.empty(len(task_ad), dtype=RESULT_DTYPE))

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

    # ---- resident arm: inputs staged once, timed region = kernels (+ the host planning of the task path)
    eng.select_slot(0); eng.stage_windows(sbuf, wlen, wl['start_ids'])
    eng.select_slot(1); eng.stage_windows(ebuf, wlen, wl['end_ids'])
    eng.select_slot(2); eng.stage_tasks(blob, seg_off, seg_len, task_ad)

    def resident_step():
        eng.select_slot(0); eng.run_windows_resident()
        eng.select_slot(1); eng.run_windows_resident()
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

    # ---- end-to-end arm: the host-buffer C-ABI calls a user makes; H2D of the inputs and D2H of the results every step
    def e2e_step():
        # the reads are all in host memory before the first alignment (porechop_abi.py:61-71): announce the blob, so
        # that it crosses PCIe under the window kernels; every byte still travels inside the timed region, every step
        eng._check(w.C_LIB.pcb_upload_reads(eng._h, _ptr(blob), blob.size), 'pcb_upload_reads')
        eng.select_slot(0)
        eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(sbuf), n, STRIDE, _ptr(wlen), _ptr(wl['start_ids']), len(wl['start_ids']), _ptr(out_s)), 'pcb_align_windows')
        eng.select_slot(1)
        eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(ebuf), n, STRIDE, _ptr(wlen), _ptr(wl['end_ids']), len(wl['end_ids']), _ptr(out_e)), 'pcb_align_windows')
        eng.select_slot(2)
        eng._check(w.C_LIB.pcb_align_tasks(eng._h, None, blob.size, _ptr(seg_off), _ptr(seg_len), _ptr(task_ad), len(task_ad), _ptr(out_m)), 'pcb_align_tasks')
        eng.wait()                       # submit / collect: the three calls above only queue; results are in the pinned buffers now

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
    eng.set_deferred(False)
    sampler.stop_flag = True; sampler.join(timeout=3)
    e2e_s = max_over_ranks(e2e_s)
    checksum = int(out_s['raw_score'].astype(np.int64).sum() + out_e['raw_score'].astype(np.int64).sum() + out_m['raw_score'].astype(np.int64).sum())

    # ---- dominant kernel alone: k_fast<TRACE, NR=24, EXACT> = the end windows against the 24-base end sequences (one launch;
    #      55% of the step in the ncu launch list), CUDA events on the library's own stream
    ad_len = np.array([len(a) for a in wl['adapters']], dtype=np.int64)
    ids24 = np.array([i for i in wl['end_ids'] if ad_len[i] == 24], dtype=np.int32)
    eng.select_slot(3); eng.stage_windows(ebuf, wlen, ids24)
    eng.run_windows_resident(); eng.synchronize()
    reps = 5
    c_k0 = eng.counters()
    eng.timer_start()
    for _ in range(reps):
        eng.run_windows_resident()
    k_ms = eng.timer_stop() / reps
    k_launches = (eng.counters()['dp_kernel_launches'] - c_k0['dp_kernel_launches']) // reps
    k_cells = int(wlen.astype(np.int64).sum()) * int(ad_len[ids24].sum())
    k_tasks = n * len(ids24)
    # algorithmic bytes of that launch: window bases read once + adapters + one 32-byte record per pair
    k_bytes = int(wlen.astype(np.int64).sum()) + int(ad_len[ids24].sum()) + 32 * k_tasks
    # DRAM traffic of the same kernel from the ncu --set full capture (profiles/kernel_r1y.md: 1.251 GB read + 3.942 GB
    # written for 32768 windows x 109 adapters = 1454 B per pair, almost all of it the checkpoint ring), per launch
    k_traffic = int(round((1.251407e9 + 3.941954e9) / (32768 * 109) * k_tasks))
    # the scan kernel of the full-length search (k_fast<SCAN>), timed with its planner / merge / window-trace launches
    eng.select_slot(2)
    eng.timer_start()
    for _ in range(reps):
        eng.run_tasks_resident()
    t_ms = eng.timer_stop() / reps
    t_cells = int((off_lens := (wl['off'][1:] - wl['off'][:-1])).astype(np.int64).sum()) * int(ad_len[wl['mid_ids']].sum())

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
        # integer-issue model: 148 SMs x 4 sub-partitions x 32 lanes x 1 instr/clk, 2 cells per lane-op
        # integer-issue roofline: a DP cell pair (two tasks packed in one register) needs at least 4
        # integer warp instructions (diagonal add, H, V, max3); profiles/pipes_r1e.txt measures 1.8 clk per integer warp
        # instruction and sub-partition at best (ALU + FMA pipes alternating) -> 64 cells per 7.2 clk per sub-partition
        instr_per_pair, clk_per_instr = 4.0, 1.8
        int_peak_gcups = 148 * 4 * 64 / (instr_per_pair * clk_per_instr) * sm_mhz * 1e6 / 1e9
        line = {
            'metric': METRIC, 'value': world * n / (ms_step / 1000.0), 'unit': 'reads/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_step, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None, 'dtype': 'int16x2', 'data': 'synthetic',
            'config': bench_config(args, wl),
            'host_binding': ('rank bound to the %d CPUs next to its GPU (NVML affinity)' % len(numa_cpus)) if numa_cpus else 'none',
            'gcups': world * wl['cells'] / (ms_step / 1000.0) / 1e9,
            'e2e': {'value': world * n * args.steps / e2e_s, 'unit': 'reads/s',
                    'h2d_bytes_per_step': int((c1['h2d_bytes'] - c0['h2d_bytes']) / args.steps),
                    'd2h_bytes_per_step': int((c1['d2h_bytes'] - c0['d2h_bytes']) / args.steps),
                    'gcups': world * wl['cells'] * args.steps / e2e_s / 1e9, 'checksum': checksum},
            'gpu_launches': int(cnt['kernel_launches']),
            'clocks': clocks,
            'roofline': {'bound': 'hbm', 'kernel': 'k_fast<TRACE,24,EXACT,16> (end windows x %d 24-base end sequences, %d launch)' % (len(ids24), k_launches), 'achieved': k_bytes / (k_ms / 1000.0) / 1e9,
                         'peak': hbm_peak, 'unit': 'GB/s', 'frac': k_bytes / (k_ms / 1000.0) / 1e9 / hbm_peak, 'traffic': k_traffic,
                         'traffic_source': 'ncu dram__bytes_read.sum + dram__bytes_write.sum per (window, adapter) pair of profiles/kernel_r1y.md x the pairs of this launch',
                         'peak_source': peak_src, 'launch_ms': k_ms,
                         'note': 'integer-issue bound kernel: see int_dp; achieved counts the algorithmic bytes only, the traffic above them is the trace checkpoints (written once in pass 1, a sixth read back in pass 2)'},
            'int_dp': {'kernel_gcups': k_cells / (k_ms / 1000.0) / 1e9, 'model_peak_gcups': int_peak_gcups,
                       'frac': k_cells / (k_ms / 1000.0) / 1e9 / int_peak_gcups,
                       'model': '148 SM x 4 sub-partitions x 64 cells / (4 integer instr x 1.8 clk) x SM clock',
                       'scan_phase_gcups': t_cells / (t_ms / 1000.0) / 1e9, 'scan_phase_frac': t_cells / (t_ms / 1000.0) / 1e9 / int_peak_gcups,
                       'sm_mhz': sm_mhz},
        }
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
