"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same seeded inputs.
Bit-exact bar: every integer field; the printed string (identities via "%f") must be identical too."""
import os
import random

import numpy as np
import pytest

from oracle_lib import oracle_align, oracle_string, random_case, rec_matches_oracle

pytestmark = pytest.mark.gpu

DEFAULT = (3, -6, -5, -2)


def _check_pairs(engine, pairs, adapters, scores):
    engine.set_scoring(scores)
    engine.set_adapters(adapters)
    recs = engine.align_pairs(pairs)
    from custom_porechop_abi_b200.cpp_function_wrappers import format_result
    bad = []
    for (read, aid), rec in zip(pairs, recs):
        o = oracle_align(read, adapters[aid], scores)
        if not rec_matches_oracle(rec, o) or format_result(rec) != oracle_string(read, adapters[aid], scores):
            bad.append((read, adapters[aid], scores, format_result(rec), oracle_string(read, adapters[aid], scores)))
    assert not bad, bad[:5]


def test_single_pair_dropin_symbol():
    from custom_porechop_abi_b200.cpp_function_wrappers import adapter_alignment
    cases = [("ACGTACGTTTGACCA", "GGGGGGG"), ("ACGTACGTTTGACCA", "TTGACCATTAGC"), ("ACGTACGTTTGACCA", "CCCACGTACG"),
             ("ACGTACGTTTGACCA", "ACGTTTG"), ("A", "C"), ("A", "A"), ("ACGT", "ACGTACGTTTGACCA"), ("NNNN", "ACGT"),
             ("acgtuu--xx", "ACGTTTNNNN"), ("", "ACGT"), ("ACGT", ""), ("", "")]
    for r, a in cases:
        for s in (DEFAULT, (3, -6, -5, -5), (1, -1, -3, -2)):
            assert adapter_alignment(r, a, list(s)) == oracle_string(r, a, s), (r, a, s)


@pytest.mark.parametrize('seed', [1, 2, 3])
def test_random_pairs_default_scheme(engine, seed):
    rng = random.Random(seed)
    adapters, pairs = [], []
    for _ in range(40):
        r, a = random_case(rng)
        adapters.append(a)
    for _ in range(3000):
        aid = rng.randrange(len(adapters))
        r, _ = random_case(rng)
        if rng.random() < 0.5:
            from oracle_lib import mutate
            a = adapters[aid]
            p = rng.randrange(0, max(1, len(r)))
            r = (r[:p] + mutate(rng, a, rng.choice([0, 0.1, 0.2])) + r[p:])[:max(len(r), 1)]
        pairs.append((r, aid))
    _check_pairs(engine, pairs, adapters, DEFAULT)


@pytest.mark.parametrize('scores', [(3, -6, -5, -5), (1, -1, -1, -1), (2, -3, -4, -1), (5, -4, -10, -1), (1, -1, -3, -2),
                                    (3, -6, -2, -5), (1, 0, 0, 0), (1, -1, 1, 1), (0, 0, 0, 0), (2, 2, -1, -2)])
def test_random_pairs_other_schemes(engine, scores):
    rng = random.Random(hash(scores) & 0xffff)
    adapters = [random_case(rng)[1] for _ in range(12)]
    pairs = []
    for _ in range(600):
        r, _ = random_case(rng, m_choices=(1, 2, 3, 7, 20, 60, 150))
        pairs.append((r, rng.randrange(len(adapters))))
    _check_pairs(engine, pairs, adapters, scores)


def test_long_adapters_multi_stripe(engine):
    rng = random.Random(5)
    adapters = [''.join(rng.choice('ACGT') for _ in range(n)) for n in (31, 32, 33, 63, 64, 65, 96, 111, 130)]
    pairs = []
    from oracle_lib import mutate
    for _ in range(400):
        aid = rng.randrange(len(adapters))
        m = rng.choice([1, 5, 40, 150, 150, 400])
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        if rng.random() < 0.7:
            p = rng.randrange(-len(adapters[aid]), m)
            ins = mutate(rng, adapters[aid], rng.choice([0, 0.05, 0.2]))
            r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
            r = r or 'A'
        pairs.append((r, aid))
    _check_pairs(engine, pairs, adapters, DEFAULT)


def _with_adapter(rng, adapter, m, rates=(0, 0.05, 0.2)):
    from oracle_lib import mutate
    r = ''.join(rng.choice('ACGT') for _ in range(m))
    if rng.random() < 0.75:
        p = rng.randrange(-len(adapter), m)
        ins = mutate(rng, adapter, rng.choice(rates))
        r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
    return r or 'A'


@pytest.mark.parametrize('scores', [DEFAULT, (2, -3, -4, -1), (1, -1, -3, -2), (5, -4, -6, -1), (4, -3, -7, -3)])
def test_two_stripe_class_pairs(engine, scores):
    """Adapters of 73..127 bases (the rapid full-sequence barcodes, adapters.py:479-498) take the two-stripe kernels;
    72 and 128 are their neighbours on either side.  Paths that end in the read's last column, in the last adapter row,
    cross the stripe boundary, start at column 0, and segments up to the widest the trace takes."""
    rng = random.Random(73)
    adapters = [''.join(rng.choice('ACGT') for _ in range(n)) for n in (72, 73, 74, 80, 96, 102, 111, 112, 113, 126, 127, 128)]
    pairs = []
    for _ in range(700):
        aid = rng.randrange(len(adapters))
        m = rng.choice([1, 2, 3, 7, 8, 9, 16, 17, 40, 71, 72, 73, 150, 150, 150, 400, 999, 1000])
        pairs.append((_with_adapter(rng, adapters[aid], m), aid))
    # the adapter itself, prefixes and suffixes of it, and reads the adapter hangs over on both sides
    for aid, a in enumerate(adapters):
        pairs += [(a, aid), (a[:50], aid), (a[-50:], aid), (a[30:-30] or 'A', aid), ('ACGT' + a + 'TTGA', aid), (a[5:] + 'GGGTTT', aid)]
    _check_pairs(engine, pairs, adapters, scores)


def test_two_stripe_class_windows_and_long_reads(engine):
    rng = random.Random(111)
    adapters = [''.join(rng.choice('ACGT') for _ in range(n)) for n in (102, 111, 24, 90, 127, 73)]
    windows = []
    for _ in range(300):
        m = rng.choice([150] * 6 + [0, 1, 17, 149])
        windows.append(_with_adapter(rng, rng.choice(adapters), m, rates=(0.05, 0.1)) if m else '')
    engine.set_scoring(DEFAULT)
    engine.set_adapters(adapters)
    buf, lens = engine.pack_windows(windows, stride=160)
    ids = [0, 1, 2, 4, 5, 3]
    out = engine.align_windows(buf, lens, ids)
    for w, r in enumerate(windows):
        for k, aid in enumerate(ids):
            assert rec_matches_oracle(out[w, k], oracle_align(r, adapters[aid], DEFAULT)), (r, adapters[aid])
    # long reads: scanned in chunks, then traced on a window around the best end cell
    pairs = []
    for _ in range(50):
        m = rng.choice([1001, 1100, 2048, 2049, 5000, 12000, 30000])
        aid = rng.randrange(len(adapters))
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        for _ in range(rng.choice([0, 1, 1, 2])):
            from oracle_lib import mutate
            p = rng.randrange(-40, m)
            ins = mutate(rng, adapters[aid], rng.choice([0, 0.05, 0.15]))
            r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
        pairs.append((r, aid))
    _check_pairs(engine, pairs, adapters, DEFAULT)


def test_fast_classes_agree_with_the_general_kernels_at_size(engine):
    """Cross-path agreement where the oracle would take minutes: 6,000 ragged windows x adapters of every kernel class
    (24 exact, < 24, 32, 40, 52, 72 rows, both two-stripe classes) through the fast kernels and, with PCB_NO_FAST, through the
    general kernels -- two independent implementations of the trace; every record identical."""
    rng = np.random.default_rng(77)
    letters = np.frombuffer(b'ACGT', dtype=np.uint8)
    adapters = [letters[rng.integers(0, 4, k)].tobytes().decode() for k in (24, 24, 17, 30, 38, 50, 68, 72, 73, 102, 111, 113, 127)]
    n, stride = 6000, 160
    lens = rng.choice([150] * 12 + [0, 1, 33, 149], n).astype(np.int32)
    buf = letters[rng.integers(0, 4, (n, stride))].astype(np.uint8)
    for w in range(0, n, 3):                      # every third window carries (most of) one of the adapters somewhere
        a = np.frombuffer(adapters[int(rng.integers(0, len(adapters)))].encode(), dtype=np.uint8)
        cut = int(rng.integers(0, len(a) // 3 + 1))
        p = int(rng.integers(0, 60))
        piece = a[cut:][:stride - p]
        buf[w, p:p + len(piece)] = piece
    engine.set_scoring(DEFAULT)
    ids = list(range(len(adapters)))
    try:
        engine.set_adapters(adapters)
        fast = engine.align_windows(buf, lens, ids)
        os.environ['PCB_NO_FAST'] = '1'
        engine.set_adapters(adapters)             # (the kernel class of an adapter is decided when the id list is split)
        general = engine.align_windows(buf, lens, ids)
    finally:
        os.environ.pop('PCB_NO_FAST', None)
        engine.set_adapters(adapters)
    assert (fast['raw_score'] > 60).sum() > 1000          # real hits among them
    bad = np.argwhere(fast != general)
    assert len(bad) == 0, (bad[:5], fast[tuple(bad[0])], general[tuple(bad[0])])


def test_full_length_search_fast_against_general_kernels(engine):
    """The same for the explicit-task path: 1,500 reads of 1 base to 60 kb x adapters of five kernel classes -- direct
    traces, chunked scans, the merge and the window traces -- on the fast kernels and, with PCB_NO_FAST, on the general ones."""
    from custom_porechop_abi_b200 import synthetic
    rng = np.random.default_rng(78)
    letters = np.frombuffer(b'ACGT', dtype=np.uint8)
    adapters = [letters[rng.integers(0, 4, k)].tobytes() for k in (24, 28, 45, 68, 102)]
    lengths = np.clip(rng.lognormal(8.0, 1.2, 1500), 1, 60000).astype(np.int64)
    blob, off = synthetic.make_reads(len(lengths), adapters, [0, 1, 2], [3, 4], seed=31, lengths=lengths, middle_rate=0.5)
    ids = np.arange(len(adapters), dtype=np.int32)
    seg_off = np.repeat(off[:-1], len(ids)); seg_len = np.repeat(lengths.astype(np.int32), len(ids)); ad = np.tile(ids, len(lengths))
    engine.set_scoring(DEFAULT)
    try:
        engine.set_adapters(adapters)
        fast = engine.align_tasks(blob, seg_off, seg_len, ad)
        os.environ['PCB_NO_FAST'] = '1'
        engine.set_adapters(adapters)
        general = engine.align_tasks(blob, seg_off, seg_len, ad)
    finally:
        os.environ.pop('PCB_NO_FAST', None)
        engine.set_adapters(adapters)
    assert (fast['raw_score'] > 60).sum() > 200
    bad = np.flatnonzero(fast != general)
    assert len(bad) == 0, (bad[:5], fast[bad[0]], general[bad[0]], seg_len[bad[0]], ad[bad[0]])


def test_windows_cross_product(engine):
    rng = random.Random(9)
    adapters = [''.join(rng.choice('ACGT') for _ in range(rng.choice([22, 24, 28, 30, 50]))) for _ in range(9)]
    from oracle_lib import mutate
    windows = []
    for _ in range(333):
        m = rng.choice([150] * 6 + [0, 1, 17, 149])
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        if m > 60 and rng.random() < 0.6:
            a = rng.choice(adapters)
            p = rng.randrange(-10, m - 10)
            ins = mutate(rng, a, 0.1)
            r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
        windows.append(r)
    engine.set_scoring(DEFAULT)
    engine.set_adapters(adapters)
    buf, lens = engine.pack_windows(windows, stride=160)
    ids = [0, 3, 4, 8, 1]
    out = engine.align_windows(buf, lens, ids)
    assert out.shape == (len(windows), len(ids))
    for w, r in enumerate(windows):
        for k, aid in enumerate(ids):
            assert rec_matches_oracle(out[w, k], oracle_align(r, adapters[aid], DEFAULT)), (r, adapters[aid])


def test_long_reads_scan_then_window(engine):
    rng = random.Random(21)
    adapters = ['AATGTACTTCGTTCAGTTACGTATTGCT', 'GCAATACGTAACTGAACGAAGT', ''.join(rng.choice('ACGT') for _ in range(70))]
    from oracle_lib import mutate
    pairs = []
    for _ in range(60):
        m = rng.choice([1100, 2500, 5000, 12000, 30000])
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        aid = rng.randrange(len(adapters))
        k = rng.choice([0, 1, 1, 2])
        for _ in range(k):
            p = rng.randrange(-20, m)
            ins = mutate(rng, adapters[aid], rng.choice([0, 0.05, 0.15]))
            r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
        if rng.random() < 0.2:
            q = rng.randrange(0, m - 50)
            r = r[:q] + '-' * 40 + r[q + 40:]      # masked stretch, as find_middle_adapters produces (nanopore_read.py:233-234)
        pairs.append((r, aid))
    _check_pairs(engine, pairs, adapters, DEFAULT)


def test_golden_vectors_from_the_reference(engine):
    """The committed vectors were produced by the reference's own library (tests/golden/make_golden.py)."""
    import json
    import os
    from custom_porechop_abi_b200.cpp_function_wrappers import format_result
    cases = json.load(open(os.path.join(os.path.dirname(__file__), 'golden', 'adapter_alignment_golden.json')))['cases']
    by_scheme = {}
    for c in cases:
        by_scheme.setdefault(tuple(c['scores']), []).append(c)
    for scores, group in by_scheme.items():
        adapters = sorted(set(c['adapter'] for c in group))
        idx = {a: i for i, a in enumerate(adapters)}
        engine.set_scoring(scores)
        engine.set_adapters(adapters)
        recs = engine.align_pairs([(c['read'], idx[c['adapter']]) for c in group])
        for c, rec in zip(group, recs):
            assert format_result(rec) == c['expect'], (c['read'][:80], c['adapter'], scores)


def test_full_size_windows_properties(engine):
    """BASELINE-sized batch (tens of millions of pairs is too many for the oracle): size-independent checks --
    determinism across two runs, agreement of the cross-product path with the explicit-task path on a sample, and
    oracle agreement on a random sample of pairs."""
    from custom_porechop_abi_b200 import synthetic
    adapters, s_ids, e_ids = synthetic.adapter_set()
    blob, off = synthetic.make_reads(20000, adapters, s_ids, e_ids, seed=11, lengths=np.full(20000, 300))
    sbuf, ebuf, lens = synthetic.end_windows(blob, off)
    engine.set_scoring(DEFAULT)
    engine.set_adapters(adapters)
    a = engine.align_windows(sbuf, lens, s_ids)
    b = engine.align_windows(sbuf, lens, s_ids)
    assert a.tobytes() == b.tobytes()
    assert int((a['raw_score'] >= 0).sum()) == a.size and int((a['read_start'] >= 0).sum()) == a.size
    rng = random.Random(4)
    sample = [(rng.randrange(len(lens)), rng.randrange(len(s_ids))) for _ in range(1500)]
    for w, k in sample:
        assert rec_matches_oracle(a[w, k], oracle_align(bytes(sbuf[w, :lens[w]]), adapters[s_ids[k]], DEFAULT))
    pairs = [(bytes(sbuf[w, :lens[w]]), s_ids[k]) for w, k in sample[:500]]
    recs = engine.align_pairs(pairs)
    for (w, k), rec in zip(sample[:500], recs):
        assert rec.tobytes() == a[w, k].tobytes()


def test_pipelined_host_buffer_paths_equal_resident(engine):
    """The host-buffer calls overlap copies with kernels (window ranges on the copy stream; the read blob in pieces,
    each with its own plan).  Same records as the plain stage / run / fetch sequence, and as the oracle on a sample."""
    import os
    from custom_porechop_abi_b200 import synthetic
    adapters, s_ids, e_ids = synthetic.adapter_set()
    engine.set_scoring(DEFAULT)
    engine.set_adapters(adapters)
    n = 49152
    blob, off = synthetic.make_reads(n, adapters, s_ids, e_ids, seed=13, lengths=np.full(n, 220))
    sbuf, ebuf, lens = synthetic.end_windows(blob, off)
    piped = engine.align_windows(ebuf, lens, e_ids)
    engine.stage_windows(ebuf, lens, e_ids)
    engine.run_windows_resident()
    plain = engine.fetch()
    assert piped.tobytes() == plain.tobytes()
    # blob in pieces: ragged reads, some straddling the piece boundaries, long and short
    rng = np.random.default_rng(5)
    lengths = np.clip(rng.lognormal(7.5, 1.0, 3000), 1, 60000).astype(np.int64)
    blob, off = synthetic.make_reads(len(lengths), adapters, s_ids, e_ids, seed=17, lengths=lengths, middle_rate=0.2)
    ids = np.array([s_ids[0], e_ids[0], s_ids[-1]], dtype=np.int32)
    seg_off = np.repeat(off[:-1], len(ids)); seg_len = np.repeat(lengths.astype(np.int32), len(ids))
    ad = np.tile(ids, len(lengths))
    one = engine.align_tasks(blob, seg_off, seg_len, ad)
    os.environ['PCB_PIECE_MIN_BYTES'] = '4096'
    try:
        for pieces in (None, '5', '16'):            # (by default a blob of this size goes in 2 pieces)
            if pieces:
                os.environ['PCB_PLAN_PIECES'] = pieces
            pieced = engine.align_tasks(blob, seg_off, seg_len, ad)
            assert pieced.tobytes() == one.tobytes(), pieces
    finally:
        del os.environ['PCB_PIECE_MIN_BYTES']
        os.environ.pop('PCB_PLAN_PIECES', None)
    # announced blob (pcb_upload_reads): it travels under the window calls made in between; NULL blob in the task call
    for env in (None, '4096'):
        if env:
            os.environ['PCB_PIECE_MIN_BYTES'] = env
        try:
            engine.upload_reads(blob)
            w1 = engine.align_windows(ebuf, lens, e_ids)
            w2 = engine.align_windows(sbuf, lens, s_ids[:2])
            up = engine.align_tasks(None, seg_off, seg_len, ad)
        finally:
            os.environ.pop('PCB_PIECE_MIN_BYTES', None)
        assert up.tobytes() == one.tobytes()
        assert w1.tobytes() == plain.tobytes() and w2.shape == (n, 2)
    # an announcement with no window call in between, then trimmed segments of the same blob (phase C after phase B)
    os.environ['PCB_PIECE_MIN_BYTES'] = '4096'
    try:
        engine.upload_reads(blob)
        cut = np.minimum(seg_len // 3, 70).astype(np.int32)
        up = engine.align_tasks(None, seg_off + cut, seg_len - 2 * cut, ad)
    finally:
        del os.environ['PCB_PIECE_MIN_BYTES']
    assert up.tobytes() == engine.align_tasks(blob, seg_off + cut, seg_len - 2 * cut, ad).tobytes()
    with pytest.raises(Exception):
        engine.align_tasks(None, seg_off, seg_len, ad)          # nothing announced any more
    # submit / collect (pcb_set_deferred + pcb_wait): the calls only queue, each in its own slot; same records
    for env in (None, '4096'):
        if env:
            os.environ['PCB_PIECE_MIN_BYTES'] = env
        try:
            with engine.deferred():
                engine.upload_reads(blob)
                engine.select_slot(0); d1 = engine.align_windows(ebuf, lens, e_ids)
                engine.select_slot(1); d2 = engine.align_windows(sbuf, lens, s_ids)
                engine.select_slot(2); d3 = engine.align_tasks(None, seg_off, seg_len, ad)
            with engine.deferred():                             # a batch that goes through in parts: three task calls on one announced blob
                engine.upload_reads(blob)
                engine.select_slot(0); engine.align_windows(ebuf, lens, e_ids)
                third = (len(ad) // 3 // len(ids)) * len(ids)
                parts = []
                for q, (a0, a1) in enumerate(((0, third), (third, 2 * third), (2 * third, len(ad)))):
                    engine.select_slot(3 + q)
                    parts.append(engine.align_tasks(None, seg_off[a0:a1], seg_len[a0:a1], ad[a0:a1]))
            assert np.concatenate(parts).tobytes() == one.tobytes()
            with engine.deferred():                             # a second batch reuses the slots and the plan buffers
                engine.select_slot(1); d4 = engine.align_windows(ebuf, lens, e_ids)
                engine.select_slot(2); d5 = engine.align_tasks(blob, seg_off + cut, seg_len - 2 * cut, ad)
        finally:
            os.environ.pop('PCB_PIECE_MIN_BYTES', None)
            engine.select_slot(0)
        assert d1.tobytes() == plain.tobytes() and d4.tobytes() == plain.tobytes()
        assert d3.tobytes() == one.tobytes() and d5.tobytes() == up.tobytes()
        assert d2.tobytes() == engine.align_windows(sbuf, lens, s_ids).tobytes()
    rs = random.Random(8)
    for t in rs.sample(range(len(ad)), 300):
        r = bytes(blob[seg_off[t]:seg_off[t] + seg_len[t]])
        assert rec_matches_oracle(pieced[t], oracle_align(r, adapters[ad[t]], DEFAULT)), t


def test_ultra_long_reads_and_barcode_sets(engine):
    """BASELINE configs[3] and [4] in miniature: 50-200 kb reads through the chunked scan (dozens of chunks per read,
    hits planted near chunk boundaries and at both read ends), and a 96-barcode set (forward + reverse, 24-mers with
    flanks of 20-40 bases -> kernel classes 2-4) against end windows, all against the oracle."""
    from oracle_lib import mutate
    rng = random.Random(31)
    ad = ['AATGTACTTCGTTCAGTTACGTATTGCT', 'GCAATACGTAACTGAACGAAGT', 'ACTTGCCTGTCGCTCTATCTTC',
          'GGCGTCTGCTTGGGTGTTTAACCTTTTTTTTTTAATGTACTTCGTTCAGTTACGTATTGCT'[:50]]
    pairs = []
    for m in (50000, 81234, 120000, 200000):
        r = [rng.choice('ACGT') for _ in range(m)]
        for aid, a in enumerate(ad):
            for p in (0, 4096 - 10, 8192 + 3, m // 2, m - len(a) - 1, m - 9):
                if rng.random() < 0.5:
                    ins = mutate(rng, a, rng.choice([0, 0.05, 0.12]))
                    r[p:p + len(ins)] = list(ins)[:max(0, m - p)]
        r = ''.join(r)[:m]
        for aid in range(len(ad)):
            pairs.append((r, aid))
    _check_pairs(engine, pairs, ad, DEFAULT)
    # barcodes
    bar = [''.join(rng.choice('ACGT') for _ in range(24)) for _ in range(48)]
    comp = {'A': 'T', 'C': 'G', 'G': 'C', 'T': 'A'}
    seqs = bar + [''.join(comp[c] for c in reversed(b)) for b in bar]
    seqs += [''.join(rng.choice('ACGT') for _ in range(k)) + b + ''.join(rng.choice('ACGT') for _ in range(8)) for b, k in zip(bar[:12], (4, 8, 12, 16, 20, 4, 8, 12, 16, 20, 3, 9))]
    windows = []
    for _ in range(400):
        w = ''.join(rng.choice('ACGT') for _ in range(150))
        b = rng.choice(seqs)
        p = rng.randrange(0, 110)
        w = (w[:p] + mutate(rng, b, rng.choice([0, 0.08, 0.2])) + w[p:])[:150]
        windows.append(w)
    engine.set_scoring(DEFAULT)
    engine.set_adapters(seqs)
    buf, lens = engine.pack_windows(windows, stride=160)
    out = engine.align_windows(buf, lens, list(range(len(seqs))))
    for w in range(0, len(windows), 3):
        for k in range(len(seqs)):
            assert rec_matches_oracle(out[w, k], oracle_align(windows[w], seqs[k], DEFAULT)), (windows[w], seqs[k])


def test_trace_queue_many_units_per_warp_and_long_paths():
    """The fast trace per batch of units: pass 1 leaves checkpoints and end cells in scratch, pass 2 sorts the tracebacks
    (first blocks by width, later blocks by rows) and runs them 64 of a kind per warp (pcb_fast.cu: k_rounds).
    A context limited to two CTAs gives every warp dozens of chunks across adapter boundaries; a 1 MB scratch budget cuts
    the launch into many batches; cheap gaps make paths that cross a dozen flag blocks (items that stay with their warp for
    many rounds of generation 1); adapters of several classes alternate.  Every record against the oracle."""
    import os
    from custom_porechop_abi_b200.engine import AlignmentEngine
    rng = random.Random(77)
    ads = [''.join(rng.choice('ACGT') for _ in range(n)) for n in (24, 24, 19, 30, 24, 37, 24, 50)]
    wins = []
    for w in range(1536):
        a = rng.choice(ads)
        kind = rng.randrange(4)
        if kind == 0:       # the two halves of an adapter far apart: one long path when gaps are cheap
            gap = ''.join(rng.choice('ACGT') for _ in range(rng.randrange(40, 115)))
            s = a[:len(a) // 2] + gap + a[len(a) // 2:]
        elif kind == 1:     # adapter overhanging either window end
            s = a[rng.randrange(1, len(a)):] + ''.join(rng.choice('ACGT') for _ in range(rng.randrange(60, 130)))
            if rng.random() < 0.5:
                s = s[::-1]
        elif kind == 2:
            s = ''.join(rng.choice('ACGT') for _ in range(rng.randrange(1, 150)))
        else:
            pos = rng.randrange(0, 100)
            s = ''.join(rng.choice('ACGT') for _ in range(pos)) + a + ''.join(rng.choice('ACGTN') for _ in range(rng.randrange(0, 40)))
        wins.append(s[:150].encode())
    os.environ['PCB_GRID_LIMIT'] = '2'
    os.environ['PCB_SCRATCH_MB'] = '1'
    try:
        eng = AlignmentEngine(0)
    finally:
        del os.environ['PCB_GRID_LIMIT']
        del os.environ['PCB_SCRATCH_MB']
    try:
        eng.set_adapters(ads)
        buf, lens = eng.pack_windows(wins, stride=160)
        ids = np.arange(len(ads), dtype=np.int32)
        for scoring in ((20, -4, -2, -1), DEFAULT):
            eng.set_scoring(scoring)
            out = eng.align_windows(buf, lens, ids)
            for w in range(len(wins)):
                for k in range(len(ads)):
                    assert rec_matches_oracle(out[w, k], oracle_align(wins[w], ads[k].encode(), scoring)), (scoring, w, k)
            # explicit tasks through the same kernels, adapters alternating from unit to unit
            blob = b''.join(wins[:512])
            off = np.concatenate([[0], np.cumsum([len(x) for x in wins[:512]])]).astype(np.int64)
            seg_off = np.repeat(off[:-1], len(ads)); seg_len = np.repeat(np.diff(off).astype(np.int32), len(ads))
            ad = np.tile(ids, 512)
            t = eng.align_tasks(blob, seg_off, seg_len, ad)
            assert t.tobytes() == out[:512].reshape(-1).tobytes()
    finally:
        eng.close()


def _ref_hits(rec, thr, at_end, end_size, min_trim):
    """The pairs nanopore_read.find_start_trim / find_end_trim keep (nanopore_read.py:187-188, :209-210), from full records."""
    from custom_porechop_abi_b200.porechop_batch import _parsed_identities, _read_ends
    partial, _ = _parsed_identities(rec)
    rs, re = rec['read_start'], _read_ends(rec)
    with np.errstate(invalid='ignore'):
        edge = (rs != 0) if at_end else (re != end_size)
        hit = (partial > thr) & edge & (re - rs >= min_trim)
    return hit


@pytest.mark.parametrize('thr,at_end,min_trim', [(75.0, False, 4), (75.0, True, 4), (60.5, False, 1), (99.999999, True, 6), (-1.0, False, 0)])
def test_reduced_outputs_match_full_records(engine, thr, at_end, min_trim):
    """pcb_align_windows_best / pcb_align_windows_hits against the same reductions done on the host from the full records
    (which the other tests check against the oracle)."""
    from custom_porechop_abi_b200 import synthetic
    from custom_porechop_abi_b200.porechop_batch import _percent_table
    engine.set_scoring(DEFAULT)
    adapters, s_ids, e_ids = synthetic.adapter_set()
    blob, off = synthetic.make_reads(700, adapters, s_ids, e_ids, seed=11, lengths=np.r_[np.full(650, 900), np.arange(0, 50) * 7])
    sbuf, ebuf, lens = synthetic.end_windows(blob, off, 150, 160)
    buf = ebuf if at_end else sbuf
    ids = np.array((e_ids if at_end else s_ids)[:57], dtype=np.int32)
    engine.set_adapters(adapters)
    rec = engine.align_windows(buf, lens, ids)
    # phase A
    best = engine.align_windows_best(buf, lens, ids)
    full = _percent_table(rec['matches'], rec['full_len']); full[rec['read_start'] == -1] = 0.0
    got = _percent_table(best['matches'], np.maximum(best['full_len'], 1)); got[best['full_len'] == 0] = 0.0
    assert np.array_equal(got, full.max(axis=0))
    # phase B
    want = np.zeros(len(ids), dtype=np.uint8); want[::3] = 1
    res = engine.align_windows_hits(buf, lens, ids, thr, at_end, 150, min_trim, want_ident=want)
    hit = _ref_hits(rec, thr, at_end, 150, min_trim)
    ii, kk = np.nonzero(hit)
    h = res.hits
    assert res.n_hits == len(ii)
    assert np.array_equal(h['window'], ii) and np.array_equal(h['id_pos'], kk)
    for f in ('read_start', 'read_end', 'raw_score', 'matches', 'aligned_len', 'full_len'):
        assert np.array_equal(h[f], rec[f][ii, kk]), f
    m, length = res.identities()
    cols = np.nonzero(want)[0]
    failed = rec['read_start'][:, cols] == -1
    assert np.array_equal(m, np.where(failed, 0, rec['matches'][:, cols])) and np.array_equal(length, np.where(failed, 0, rec['full_len'][:, cols]))
    # barcode binning input: the two best identities per window, descending, the first in adapter order among equals
    top = engine.fetch_top2(len(lens))
    sc = np.where(length > 0, m / np.maximum(length, 1), 0.0)              # exact enough to order ratios of lengths < 512
    order = np.argsort(-sc, axis=1, kind='stable')
    assert np.array_equal(top['best'], order[:, 0]) and np.array_equal(top['second'], order[:, 1])
    rows = np.arange(len(lens))
    assert np.array_equal(top['best_ident'], res.ident[rows, order[:, 0]]) and np.array_equal(top['second_ident'], res.ident[rows, order[:, 1]])
    # what phase C needs of phase B: the trim amount per window (find_start_trim / find_end_trim before extra_trim_size)
    trim = engine.fetch_trims(len(lens))
    value = np.where(hit, (150 - rec['read_start']) if at_end else (rec['read_end'] + 1), -1)
    assert np.array_equal(trim, value.max(axis=1))
    # the same through submit / collect
    with engine.deferred():                      # every queued call in its own slot
        engine.select_slot(1)
        res2 = engine.align_windows_hits(buf, lens, ids, thr, at_end, 150, min_trim, want_ident=want, hits_cap=rec.size)
        trim2 = engine.fetch_trims(len(lens))
        engine.select_slot(2)
        best2 = engine.align_windows_best(buf, lens, ids)
    engine.select_slot(0)
    assert res2.n_hits == res.n_hits and np.array_equal(res2.hits, res.hits) and np.array_equal(res2.ident, res.ident)
    assert np.array_equal(best2, best) and np.array_equal(trim2, trim)


def test_batch_in_parts_as_the_end_to_end_step_runs_it(engine):
    """bench.py's end-to-end step as a parity test: a read set goes through in two parts; phase B of both parts is queued
    (reduced outputs: hits, identities, trims), part by part its results are collected with pcb_wait_slot while the rest keeps
    running, the trims become that part's phase-C segments, and the two task calls share one announced blob.  Every array
    against the same calls made one by one."""
    from custom_porechop_abi_b200 import synthetic
    engine.set_scoring(DEFAULT)
    adapters, s_ids, e_ids = synthetic.adapter_set()
    lengths = np.r_[np.full(1500, 2500), np.arange(1, 301) * 9]
    blob, off = synthetic.make_reads(len(lengths), adapters, s_ids, e_ids, seed=23, lengths=lengths, middle_rate=0.2)
    sbuf, ebuf, lens = synthetic.end_windows(blob, off, 150, 160)
    n = len(lens)
    engine.set_adapters(adapters)
    s_ids, e_ids = np.array(s_ids[:40], dtype=np.int32), np.array(e_ids[:40], dtype=np.int32)
    mids = np.array([s_ids[0], e_ids[0], s_ids[5]], dtype=np.int32)
    want = np.zeros(40, dtype=np.uint8); want[::4] = 1
    cut = [0, n // 2, n]
    rlen = np.diff(off)

    def segments(a, b, trim_s, trim_e):
        st = np.where(trim_s >= 0, trim_s.astype(np.int64) + 2, 0)
        et = np.where(trim_e >= 0, trim_e.astype(np.int64) + 2, 0)
        lo = np.minimum(st, rlen[a:b]); hi = np.maximum(rlen[a:b] - et, lo)
        return np.repeat(off[a:b] + lo, len(mids)), np.repeat(hi - lo, len(mids)).astype(np.int32), np.tile(mids, b - a)

    # one by one
    ref = []
    for p in range(2):
        a, b = cut[p], cut[p + 1]
        hs = engine.align_windows_hits(sbuf[a:b], lens[a:b], s_ids, 75.0, False, 150, 4, want_ident=want)
        ts = engine.fetch_trims(b - a)
        he = engine.align_windows_hits(ebuf[a:b], lens[a:b], e_ids, 75.0, True, 150, 4, want_ident=want)
        te = engine.fetch_trims(b - a)
        so, sl, sa = segments(a, b, ts, te)
        ref.append((hs.hits.copy(), hs.ident.copy(), ts, he.hits.copy(), he.ident.copy(), te, engine.align_tasks(blob, so, sl, sa)))
    assert sum(len(r[0]) + len(r[3]) for r in ref) > 500 and any((r[2] >= 0).any() for r in ref)
    # queued, as the step does it
    os.environ['PCB_PIECE_MIN_BYTES'] = '4096'            # the blob travels in 16 pieces, paced by the window calls
    try:
        engine.set_deferred(True)
        engine.upload_reads(blob)
        q = []
        for p in range(2):
            a, b = cut[p], cut[p + 1]
            engine.select_slot(2 * p)
            hs = engine.align_windows_hits(sbuf[a:b], lens[a:b], s_ids, 75.0, False, 150, 4, want_ident=want, hits_cap=40 * (b - a))
            ts = engine.fetch_trims(b - a)
            engine.select_slot(2 * p + 1)
            he = engine.align_windows_hits(ebuf[a:b], lens[a:b], e_ids, 75.0, True, 150, 4, want_ident=want, hits_cap=40 * (b - a))
            te = engine.fetch_trims(b - a)
            q.append((hs, ts, he, te))
        outs = []
        for p in range(2):
            a, b = cut[p], cut[p + 1]
            hs, ts, he, te = q[p]
            engine.wait_slot(2 * p); engine.wait_slot(2 * p + 1)          # this part is on the host, the other still runs
            assert np.array_equal(hs.hits, ref[p][0]) and np.array_equal(hs.ident, ref[p][1]) and np.array_equal(ts, ref[p][2]), p
            assert np.array_equal(he.hits, ref[p][3]) and np.array_equal(he.ident, ref[p][4]) and np.array_equal(te, ref[p][5]), p
            so, sl, sa = segments(a, b, ts, te)
            engine.select_slot(4 + p)
            outs.append(engine.align_tasks(None, so, sl, sa))
        engine.wait()
    finally:
        os.environ.pop('PCB_PIECE_MIN_BYTES', None)
        engine.set_deferred(False)
        engine.select_slot(0)
    for p in range(2):
        assert outs[p].tobytes() == ref[p][6].tobytes(), p


@pytest.mark.parametrize('scores', [(7, -6, -5, -1), (10, -10, -10, -1), (4, -3, -7, -3)])
def test_user_schemes_on_kilobase_segments(engine, scores):
    """--scoring_scheme values for which the trace window of a long read (warm-up + span) is wider than
    the fast trace kernels take, or that the fast kernels do not take at all: ordinary 1-3 kb reads must still align (direct
    tasks up to the fast limit, then scan + window, the window on the general kernel where it is too wide)."""
    rng = random.Random(sum(scores) & 0xff)
    ads = [''.join(rng.choice('ACGT') for _ in range(n)) for n in (24, 52, 33)]
    pairs = []
    for L in (900, 1001, 1400, 1790, 2600, 3000):
        for aid, a in enumerate(ads):
            r = ''.join(rng.choice('ACGT') for _ in range(L))
            p = rng.randrange(100, L - 100)
            from oracle_lib import mutate
            r = r[:p] + mutate(rng, a, 0.1) + r[p + len(a):]
            pairs.append((r, aid))
    _check_pairs(engine, pairs, ads, scores)


def test_windows_cut_from_the_read_blob_on_the_device(engine):
    """pcb_stage_windows_from_reads: seq[:end_size] / seq[-end_size:] of every read taken out of the announced blob by a
    kernel give the same records as windows packed on the host (short and empty reads, lower case, a blob in pieces)."""
    import os
    from custom_porechop_abi_b200 import synthetic
    engine.set_scoring(DEFAULT)
    adapters, s_ids, e_ids = synthetic.adapter_set()
    lengths = np.r_[np.arange(0, 40), np.full(300, 700), [149, 150, 151, 3000]]
    blob, off = synthetic.make_reads(len(lengths), adapters, s_ids, e_ids, seed=21, lengths=lengths)
    blob = blob.copy(); blob[off[50]:off[60]] += 32           # lower case reads
    sbuf, ebuf, wlen = synthetic.end_windows(blob, off, 150, 160)
    engine.set_adapters(adapters)
    ids = np.array(s_ids[:9] + e_ids[:9], dtype=np.int32)
    for piece_min in (None, '4096'):
        if piece_min:
            os.environ['PCB_PIECE_MIN_BYTES'] = piece_min
        try:
            for at_end, buf in ((False, sbuf), (True, ebuf)):
                want = engine.align_windows(buf, wlen, ids)
                engine.upload_reads(blob)
                n = engine.stage_windows_from_reads(off[:-1], np.diff(off), 150, at_end)
                best = engine.align_windows_best(None, n, ids)
                res = engine.align_windows_hits(None, n, ids, 75.0, at_end, 150, 4, hits_cap=want.size)
                engine.run_windows_resident()
                got = engine.fetch((n, len(ids)))
                assert got.tobytes() == want.tobytes()
                ref_best = engine.align_windows_best(buf, wlen, ids)
                ref_hits = engine.align_windows_hits(buf, wlen, ids, 75.0, at_end, 150, 4, hits_cap=want.size)
                assert np.array_equal(best, ref_best) and np.array_equal(res.hits, ref_hits.hits)
        finally:
            os.environ.pop('PCB_PIECE_MIN_BYTES', None)
