"""Kernel-only timing of end windows x the 68- and 63-base full-sequence barcode sets: python scripts/exp/kbench_long.py [n_reads]
(PCB_NO_FAST=1 sends them to the general kernel k_align, as before the 72-row class existed)"""
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
eng.stage_windows(sbuf, lens, ids)
eng.run_windows_resident(); eng.synchronize()
eng.timer_start()
for _ in range(3):
    eng.run_windows_resident()
ms = eng.timer_stop() / 3
cells = int(lens.sum()) * int(ad_len[ids].sum())
out = eng.fetch()
print('%d x %d-base adapters  %8.3f ms  %7.1f GCUPS  checksum %d  (%s)' % (len(ids), ad_len[ids[0]], ms, cells / ms / 1e6,
      int(out['raw_score'].astype(np.int64).sum() + out['read_end'].astype(np.int64).sum()), 'general kernel' if os.environ.get('PCB_NO_FAST') else 'fast class'))
