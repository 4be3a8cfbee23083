"""Wall time of the three batched phase functions (porechop_batch.py) on synthetic reads, host logic included:
python scripts/exp/batch_phases.py [n_reads]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from custom_porechop_abi_b200 import synthetic, porechop_batch as pb
from custom_porechop_abi_b200.engine import AlignmentEngine


class Read(object):                       # what the phases touch of the reference's NanoporeRead (nanopore_read.py:32-75)
    def __init__(self, name, seq):
        self.name, self.seq = name, seq
        self.start_trim_amount = self.end_trim_amount = 0
        self.start_adapter_alignments, self.end_adapter_alignments = [], []
        self.middle_adapter_positions, self.middle_trim_positions, self.middle_hit_str = set(), set(), ''
        self.start_barcode_scores, self.end_barcode_scores = {}, {}

    def get_seq_with_start_end_adapters_trimmed(self):
        if not self.start_trim_amount and not self.end_trim_amount:
            return self.seq
        return self.seq[self.start_trim_amount:len(self.seq) - self.end_trim_amount]

    def determine_barcode(self, *a):
        pass


class AdapterSet(object):                 # adapters.py:18-53
    def __init__(self, name, start, end):
        self.name, self.start_sequence, self.end_sequence = name, start, end
        self.best_start_score = self.best_end_score = 0.0

    def best_start_or_end_score(self):
        return max(self.best_start_score, self.best_end_score)

    def is_barcode(self):
        return False


n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
adapters, s_ids, e_ids = synthetic.adapter_set()
blob, off = synthetic.make_reads(n, adapters, s_ids, e_ids, seed=3)
text = blob.tobytes().decode()
reads = [Read('r%d' % i, text[off[i]:off[i + 1]]) for i in range(n)]
sets = []
for k in range(max(len(s_ids), len(e_ids))):
    st = ('s%d' % k, adapters[s_ids[k]].decode() if isinstance(adapters[s_ids[k]], bytes) else adapters[s_ids[k]]) if k < len(s_ids) else None
    en = ('e%d' % k, adapters[e_ids[k]].decode() if isinstance(adapters[e_ids[k]], bytes) else adapters[e_ids[k]]) if k < len(e_ids) else None
    sets.append(AdapterSet('set %d' % k, st, en))
eng = AlignmentEngine(0)
vals = (3, -6, -5, -2)
for rep in range(2):
    for r in reads:
        r.__init__(r.name, r.seq)
    for a in sets:
        a.best_start_score = a.best_end_score = 0.0
    t0 = time.perf_counter()
    matching = pb.find_matching_adapter_sets(reads[:10000], 0, 150, vals, None, 90.0, 1, sets, engine=eng)
    t1 = time.perf_counter()
    pb.find_adapters_at_read_ends(reads, matching, 0, 150, 2, 75.0, vals, None, 4, 1, False, 75.0, 5.0, False, None, engine=eng)
    t2 = time.perf_counter()
    pb.find_adapters_in_read_middles(reads, matching, 0, 90.0, 10, 100, vals, None, 1, False, engine=eng)
    t3 = time.perf_counter()
    print('rep %d  n=%d  matching sets %d | phase A (10000 reads x %d sets) %.3f s | phase B %.3f s | phase C %.3f s | trimmed starts %d ends %d middles %d' % (
        rep, n, len(matching), len(sets), t1 - t0, t2 - t1, t3 - t2, sum(r.start_trim_amount > 0 for r in reads),
        sum(r.end_trim_amount > 0 for r in reads), sum(bool(r.middle_adapter_positions) for r in reads)), flush=True)
# the same with the reads as a ReadBlob of a FASTQ file (no Python string handling per phase)
import tempfile
from custom_porechop_abi_b200.read_blob import load_read_blob
with tempfile.TemporaryDirectory() as tmp:
    path = os.path.join(tmp, 'reads.fastq')
    with open(path, 'w') as f:
        for r in reads:
            f.write('@%s\n%s\n+\n%s\n' % (r.name, r.seq, 'I' * len(r.seq)))
    t0 = time.perf_counter()
    rb = load_read_blob(path)
    t_load = time.perf_counter() - t0
for rep in range(2):
    for r in reads:
        r.__init__(r.name, r.seq)
    t1 = time.perf_counter()
    pb.find_adapters_at_read_ends(reads, matching, 0, 150, 2, 75.0, vals, None, 4, 1, False, 75.0, 5.0, False, None, engine=eng, read_blob=rb)
    t2 = time.perf_counter()
    pb.find_adapters_in_read_middles(reads, matching, 0, 90.0, 10, 100, vals, None, 1, False, engine=eng, read_blob=rb)
    t3 = time.perf_counter()
    print('read_blob rep %d  load %.3f s | phase B %.3f s | phase C %.3f s | trimmed starts %d ends %d middles %d' % (
        rep, t_load, t2 - t1, t3 - t2, sum(r.start_trim_amount > 0 for r in reads), sum(r.end_trim_amount > 0 for r in reads),
        sum(bool(r.middle_adapter_positions) for r in reads)), flush=True)

if os.environ.get('PROFILE_C'):
    import cProfile, pstats
    for r in reads:
        r.middle_adapter_positions, r.middle_trim_positions, r.middle_hit_str = set(), set(), ''
    pr = cProfile.Profile(); pr.enable()
    pb.find_adapters_in_read_middles(reads, matching, 0, 90.0, 10, 100, vals, None, 1, False, engine=eng)
    pr.disable()
    pstats.Stats(pr).sort_stats('tottime').print_stats(14)
if os.environ.get('PROFILE_B'):
    import cProfile, pstats
    for r in reads:
        r.__init__(r.name, r.seq)
    pr = cProfile.Profile(); pr.enable()
    pb.find_adapters_at_read_ends(reads, matching, 0, 150, 2, 75.0, vals, None, 4, 1, False, 75.0, 5.0, False, None, engine=eng, read_blob=rb)
    pr.disable()
    pstats.Stats(pr).sort_stats('tottime').print_stats(12)
