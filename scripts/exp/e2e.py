"""The bench's end-to-end step alone (submit / collect), wall clock per step.  python scripts/exp/e2e.py [n_reads] [steps]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
from custom_porechop_abi_b200.engine import AlignmentEngine, RESULT_DTYPE, _ptr
from custom_porechop_abi_b200 import cpp_function_wrappers as w
n = int(sys.argv[1]) if len(sys.argv) > 1 else 131072
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
wl = bench.workload(n, 1)
eng = AlignmentEngine(0)
eng.set_adapters(wl['adapters'])
keep = []
def P(a):
    t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory(); keep.append(t); return t.numpy()
def pinned_out(shape):
    t = torch.empty(tuple(shape) + (8,), dtype=torch.int32).pin_memory(); keep.append(t)
    return t.numpy().view(RESULT_DTYPE).reshape(shape)
sbuf, ebuf, wlen, blob = P(wl['sbuf']), P(wl['ebuf']), P(wl['wlen']), P(wl['blob'])
seg_off, seg_len, task_ad = P(wl['seg_off']), P(wl['seg_len']), P(wl['task_ad'])
sid = np.ascontiguousarray(wl['start_ids']); eid = np.ascontiguousarray(wl['end_ids'])
out_s = pinned_out((n, len(sid))); out_e = pinned_out((n, len(eid))); out_m = pinned_out((len(task_ad),))
stride = sbuf.shape[1]
def step(mark=False):
    t = [time.perf_counter()]
    eng._check(w.C_LIB.pcb_upload_reads(eng._h, _ptr(blob), blob.size), 'x')
    eng.select_slot(0)
    eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(sbuf), n, stride, _ptr(wlen), _ptr(sid), len(sid), _ptr(out_s)), 'x'); t.append(time.perf_counter())
    eng.select_slot(1)
    eng._check(w.C_LIB.pcb_align_windows(eng._h, _ptr(ebuf), n, stride, _ptr(wlen), _ptr(eid), len(eid), _ptr(out_e)), 'x'); t.append(time.perf_counter())
    eng.select_slot(2)
    eng._check(w.C_LIB.pcb_align_tasks(eng._h, None, blob.size, _ptr(seg_off), _ptr(seg_len), _ptr(task_ad), len(task_ad), _ptr(out_m)), 'x'); t.append(time.perf_counter())
    eng.wait(); t.append(time.perf_counter())
    return [(b - a) * 1e3 for a, b in zip(t[:-1], t[1:])]
for deferred in (False, True):
    eng.set_deferred(deferred)
    for _ in range(2):
        step()
    t0 = time.perf_counter()
    parts = np.array([step() for _ in range(steps)])
    ms = (time.perf_counter() - t0) / steps * 1e3
    print('deferred=%d ranges=%s  %.2f ms/step  %.0f reads/s   host ms per call: win %.2f win %.2f tasks %.2f wait %.2f' % (
        deferred, os.environ.get('PCB_RANGES', '8'), ms, n / ms * 1e3, *parts.mean(axis=0)), flush=True)
eng.set_deferred(False)
