"""One launch class of the window cross product (for ncu): python scripts/exp/one.py [n_reads]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from custom_porechop_abi_b200 import synthetic
from custom_porechop_abi_b200.engine import AlignmentEngine
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
adapters, s_ids, e_ids = synthetic.adapter_set()
blob, off = synthetic.make_reads(n, adapters, s_ids, e_ids, seed=5, lengths=np.full(n, 400))
sbuf, ebuf, lens = synthetic.end_windows(blob, off)
eng = AlignmentEngine(0)
eng.set_adapters(adapters)
ids = [i for i in s_ids if len(adapters[i]) == 24]
eng.stage_windows(sbuf, lens, ids)
for _ in range(2):
    eng.run_windows_resident()
eng.synchronize()
print('ok')
