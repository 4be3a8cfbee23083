"""Where a bench step spends its time: per-submit resident time, summed DP-kernel time, and the host-buffer calls.
python scripts/exp/phases.py [n_reads]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
from custom_porechop_abi_b200.engine import AlignmentEngine

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
wl = bench.workload(n, 1)
eng = AlignmentEngine(0)
eng.set_adapters(wl['adapters'])
eng.select_slot(0); eng.stage_windows(wl['sbuf'], wl['wlen'], wl['start_ids'])
eng.select_slot(1); eng.stage_windows(wl['ebuf'], wl['wlen'], wl['end_ids'])
eng.select_slot(2); eng.stage_tasks(wl['blob'], wl['seg_off'], wl['seg_len'], wl['task_ad'])
runs = [('start windows', 0, eng.run_windows_resident), ('end windows', 1, eng.run_windows_resident), ('full reads', 2, eng.run_tasks_resident)]
for name, slot, fn in runs:
    eng.select_slot(slot); fn(); eng.synchronize()
for timing in (0, 1):
    eng.set_kernel_timing(timing)
    for name, slot, fn in runs:
        eng.select_slot(slot)
        eng.reset_counters()
        t0 = time.perf_counter()
        eng.timer_start()
        for _ in range(3):
            fn()
        ms = eng.timer_stop() / 3
        wall = (time.perf_counter() - t0) / 3 * 1e3
        c = eng.counters()
        print('%-14s ktiming=%d  device %8.2f ms  wall %8.2f ms  dp kernels %8.2f ms in %d launches  %7.1f GCUPS' % (
            name, timing, ms, wall, c['dp_kernel_ms'] / 3, c['dp_kernel_launches'] // 3, c['dp_cells'] / 3 / ms / 1e6), flush=True)
eng.set_kernel_timing(0)
# host-buffer path split into its parts, pinned buffers (as bench.py)
import torch
def pin(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory(); return t, t.numpy()
keep = []
def P(a):
    t, v = pin(a); keep.append(t); return v
sbuf, wlen, blob = P(wl['sbuf']), P(wl['wlen']), P(wl['blob'])
seg_off, seg_len, task_ad = P(wl['seg_off']), P(wl['seg_len']), P(wl['task_ad'])
def timed(name, fn):
    eng.synchronize(); t0 = time.perf_counter(); r = fn(); eng.synchronize()
    print('   %-28s %8.2f ms' % (name, (time.perf_counter() - t0) * 1e3), flush=True); return r
for rep in range(2):
    print('rep', rep)
    eng.select_slot(0)
    timed('stage_windows (start)', lambda: eng.stage_windows(sbuf, wlen, wl['start_ids']))
    timed('run_windows', eng.run_windows_resident)
    timed('fetch (pageable dst)', eng.fetch)
    eng.select_slot(2)
    timed('stage_tasks', lambda: eng.stage_tasks(blob, seg_off, seg_len, task_ad))
    timed('run_tasks (plan + kernels)', eng.run_tasks_resident)
    timed('run_tasks (kernels)', eng.run_tasks_resident)
    timed('fetch', eng.fetch)

# the three host-buffer calls of a bench step, pinned in and out
from custom_porechop_abi_b200.engine import RESULT_DTYPE, _ptr
from custom_porechop_abi_b200 import cpp_function_wrappers as w
def pinned_out(shape):
    t = torch.empty(tuple(shape) + (8,), dtype=torch.int32).pin_memory(); keep.append(t)
    return t.numpy().view(RESULT_DTYPE).reshape(shape)
ebuf = P(wl['ebuf'])
out_s = pinned_out((n, len(wl['start_ids']))); out_e = pinned_out((n, len(wl['end_ids']))); out_m = pinned_out((len(task_ad),))
sid = np.ascontiguousarray(wl['start_ids']); eid = np.ascontiguousarray(wl['end_ids'])
for rep in range(3):
    print('e2e rep', rep)
    eng.select_slot(0)
    timed('align_windows start', lambda: eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(sbuf), n, 160, _ptr(wlen), _ptr(sid), len(sid), _ptr(out_s)), 'x'))
    eng.select_slot(1)
    timed('align_windows end', lambda: eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(ebuf), n, 160, _ptr(wlen), _ptr(eid), len(eid), _ptr(out_e)), 'x'))
    eng.select_slot(2)
    timed('align_tasks', lambda: eng._check(w.C_LIB.pcb_align_tasks(eng._h, _ptr(blob), blob.size, _ptr(seg_off), _ptr(seg_len), _ptr(task_ad), len(task_ad), _ptr(out_m)), 'x'))
def timed_ns(name, fn):
    t0 = time.perf_counter(); r = fn()
    print('   %-28s %8.2f ms (no sync)' % (name, (time.perf_counter() - t0) * 1e3), flush=True); return r
for rep in range(3):
    print('e2e announced rep', rep)
    eng.synchronize(); t00 = time.perf_counter()
    timed_ns('upload_reads', lambda: eng._check(w.C_LIB.pcb_upload_reads(eng._h, _ptr(blob), blob.size), 'x'))
    eng.select_slot(0)
    timed_ns('align_windows start', lambda: eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(sbuf), n, 160, _ptr(wlen), _ptr(sid), len(sid), _ptr(out_s)), 'x'))
    eng.select_slot(1)
    timed_ns('align_windows end', lambda: eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(ebuf), n, 160, _ptr(wlen), _ptr(eid), len(eid), _ptr(out_e)), 'x'))
    eng.select_slot(2)
    timed_ns('align_tasks (announced)', lambda: eng._check(w.C_LIB.pcb_align_tasks(eng._h, None, blob.size, _ptr(seg_off), _ptr(seg_len), _ptr(task_ad), len(task_ad), _ptr(out_m)), 'x'))
    eng.synchronize(); print('   step total %8.2f ms' % ((time.perf_counter() - t00) * 1e3), flush=True)
