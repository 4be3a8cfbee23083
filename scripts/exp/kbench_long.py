"""Kernel-only timing of end windows x full-sequence barcode sets: python scripts/exp/kbench_long.py [n_reads] [native|rapid]
native: the 68-base start sets of the barcodes workload (72-row class; PCB_NO_FAST=1 sends them to the general kernel k_align)
rapid:  12 x 102-base + 12 x 111-base start sets, adapters.py:479-498 shape (two-stripe class; PCB_NO_TWO=1 -> general kernel)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import bench
from custom_porechop_abi_b200 import synthetic
from custom_porechop_abi_b200.engine import AlignmentEngine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
adapters, s_ids, e_ids, _ = bench.barcode_adapter_set()
ad_len = np.array([len(a) for a in adapters])
blob, off = synthetic.make_reads(n, adapters, s_ids, e_ids, seed=5, lengths=np.full(n, 400))
sbuf, ebuf, lens = synthetic.end_windows(blob, off)
eng = AlignmentEngine(0)
eng.set_adapters(adapters)
ids = [i for i in s_ids if ad_len[i] > 52]
if len(sys.argv) > 2 and sys.argv[2] == 'rapid':
    rng = np.random.default_rng(11)
    rnd = lambda k: np.frombuffer(b'ACGT', dtype=np.uint8)[rng.integers(0, 4, k)].tobytes()
    head, tail, m1, m2 = rnd(21), rnd(50), rnd(7), rnd(16)
    rapid = [head + m1 + adapters[i] + tail for i in range(12)] + [head + m2 + adapters[i] + tail for i in range(12)]
    ids = list(range(len(adapters), len(adapters) + len(rapid)))
    adapters = list(adapters) + rapid
    ad_len = np.array([len(a) for a in adapters])
    # every fourth read carries one of them at its start
    rr = np.random.default_rng(12)
    for r in range(0, n, 4):
        a = np.frombuffer(rapid[int(rr.integers(0, len(rapid)))], dtype=np.uint8)
        cut = int(rr.integers(0, 30))
        sbuf[r, :len(a) - cut] = a[cut:]
    eng.set_adapters(adapters)
eng.stage_windows(sbuf, lens, ids)
eng.run_windows_resident(); eng.synchronize()
eng.timer_start()
for _ in range(3):
    eng.run_windows_resident()
ms = eng.timer_stop() / 3
cells = int(lens.sum()) * int(ad_len[ids].sum())
out = eng.fetch()
print('%d x %d-base adapters  %8.3f ms  %7.1f GCUPS  checksum %d  (%s)' % (len(ids), ad_len[ids[0]], ms, cells / ms / 1e6,
      int(out['raw_score'].astype(np.int64).sum() + out['read_end'].astype(np.int64).sum()), 'general kernel' if os.environ.get('PCB_NO_FAST') or os.environ.get('PCB_NO_TWO') else 'fast class'))
