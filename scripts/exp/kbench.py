"""Kernel-only timing of the window cross product (experiments): python scripts/exp/kbench.py [n_reads]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from custom_porechop_abi_b200 import synthetic
from custom_porechop_abi_b200.engine import AlignmentEngine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
only24 = len(sys.argv) > 2 and sys.argv[2] == '24'
adapters, s_ids, e_ids = synthetic.adapter_set()
blob, off = synthetic.make_reads(n, adapters, s_ids, e_ids, seed=5, lengths=np.full(n, 400))
sbuf, ebuf, lens = synthetic.end_windows(blob, off)
eng = AlignmentEngine(0)
eng.set_adapters(adapters)
ad_len = np.array([len(a) for a in adapters])
for name, ids in (('all start ids', s_ids), ('24-mers only', [i for i in s_ids if ad_len[i] == 24]), ('<24', [i for i in s_ids if ad_len[i] < 24]), ('25..32', [i for i in s_ids if 24 < ad_len[i] <= 32]), ('33..40', [i for i in s_ids if 32 < ad_len[i] <= 40]), ('41..52', [i for i in s_ids if 40 < ad_len[i] <= 52])):
    if only24 and name != '24-mers only':
        continue
    eng.stage_windows(sbuf, lens, ids)
    eng.run_windows_resident(); eng.synchronize()
    eng.timer_start()
    for _ in range(3):
        eng.run_windows_resident()
    ms = eng.timer_stop() / 3
    cells = int(lens.sum()) * int(ad_len[ids].sum())
    out = eng.fetch()
    print('%-16s %8.3f ms  %7.1f GCUPS  checksum %d' % (name, ms, cells / ms / 1e6, int(out['raw_score'].astype(np.int64).sum() + out['read_end'].astype(np.int64).sum())), flush=True)
