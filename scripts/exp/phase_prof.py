"""Where the time of bench.phase_e2e goes on the host: python scripts/exp/phase_prof.py [n_reads ...]  (cProfile, top functions)"""
import cProfile, io, os, pstats, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch
import bench
from custom_porechop_abi_b200.engine import AlignmentEngine
sys.argv = sys.argv[:1] + ['--no-cpu-baseline']
ns = [int(x) for x in os.environ.get('PHASE_N', '65536,131072').split(',')]
args = bench.parse_args()
wl = bench.workload(args, 0, 1)
eng = AlignmentEngine(0)
eng.set_adapters(wl['adapters'])
t = torch.from_numpy(np.ascontiguousarray(wl['blob'])).pin_memory()
blob = t.numpy()
for n in ns:
    r = bench.phase_e2e(eng, wl, blob, n_reads=n)
    print(n, '%.3f M reads/s' % (r['value'] / 1e6), {k: round(v * 1e3, 1) for k, v in r['seconds'].items()}, 'first pass %.1f ms' % (r['first_pass_seconds'] * 1e3),
          'h2d per pass %.3f GB' % (r['h2d_bytes_per_pass'] / 1e9), flush=True)
    pr = cProfile.Profile()
    pr.enable()
    bench.phase_e2e(eng, wl, blob, n_reads=n, reps=1)
    pr.disable()
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats('cumulative').print_stats(38)
    print('\n'.join(l[:170] for l in s.getvalue().split('\n')[4:60]), flush=True)
