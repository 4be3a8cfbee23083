"""The batched phase functions (custom_porechop_abi_b200/porechop_batch.py) against the state the reference's own
Python + C++ left on its reads and adapter sets (tests/golden/phases_golden.json, made by make_phase_golden.py).
CPU: the host logic over a stand-in engine that answers with the oracle.  GPU: the same through the CUDA library."""
import json
import os

import numpy as np
import pytest

from oracle_lib import oracle_align

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden', 'phases_golden.json')
OUT_GOLDEN = os.path.join(os.path.dirname(__file__), 'golden', 'output_golden.json')


class Read(object):
    """The attributes of the reference's NanoporeRead that the three phases touch (nanopore_read.py:32-75)."""

    def __init__(self, name, seq, quals=''):
        self.name = name
        self.seq = seq.upper()
        self.rna = self.seq.count('U') > self.seq.count('T')
        if self.rna:
            self.seq = self.seq.replace('U', 'T')
        self.quals = quals + '+' * max(0, len(seq) - len(quals))
        self.start_trim_amount = 0
        self.end_trim_amount = 0
        self.start_adapter_alignments = []
        self.end_adapter_alignments = []
        self.middle_adapter_positions = set()
        self.middle_trim_positions = set()
        self.middle_hit_str = ''
        self.start_barcode_scores = {}
        self.end_barcode_scores = {}

    def get_seq_with_start_end_adapters_trimmed(self):
        if not self.start_trim_amount and not self.end_trim_amount:
            return self.seq
        return self.seq[self.start_trim_amount:len(self.seq) - self.end_trim_amount]

    barcode_call = 'none'
    albacore_barcode_call = None

    def determine_barcode(self, barcode_threshold, barcode_diff, require_two_barcodes):
        from custom_porechop_abi_b200.porechop_batch import determine_barcode
        determine_barcode(self, barcode_threshold, barcode_diff, require_two_barcodes)


class AdapterSet(object):
    """The attributes of the reference's Adapter (adapters.py:18-53) the phases touch."""

    def __init__(self, d):
        self.name = d['name']
        self.start_sequence = tuple(d['start_sequence']) if d['start_sequence'] else None
        self.end_sequence = tuple(d['end_sequence']) if d['end_sequence'] else None
        self.best_start_score, self.best_end_score = 0.0, 0.0

    def best_start_or_end_score(self):
        return max(self.best_start_score, self.best_end_score)

    def is_barcode(self):
        return self.name.startswith('Barcode ')

    def barcode_direction(self):
        return 'reverse' if '_rev' in self.start_sequence[0] else 'forward'

    def get_barcode_name(self):
        names = [self.name] + ([self.start_sequence[0]] if self.start_sequence else []) + \
            ([self.end_sequence[0]] if self.end_sequence else [])
        return sorted(names, key=lambda x: len(x))[0].replace(' ', '_')


class OracleEngine(object):
    """Same calls as AlignmentEngine, answered by the CPU oracle (test infrastructure: checks the host logic only)."""

    def __init__(self):
        from custom_porechop_abi_b200.engine import RESULT_DTYPE
        self.dtype = RESULT_DTYPE
        self.scoring = (3, -6, -5, -2)
        self.adapters = []

    def set_scoring(self, vals):
        self.scoring = tuple(int(v) for v in vals)

    def set_adapters(self, seqs):
        self.adapters = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]

    def _one(self, seq, aid):
        o = oracle_align(seq, self.adapters[aid], self.scoring)
        return (o.read_start, o.read_end, o.adapter_start, o.adapter_end, o.raw_score, o.full_matches, o.aligned_len, o.full_len)

    def align_windows(self, buf, lens, ids):
        out = np.empty((len(lens), len(ids)), dtype=self.dtype)
        for i in range(len(lens)):
            w = bytes(buf[i, :lens[i]])
            for k, aid in enumerate(ids):
                out[i, k] = self._one(w, int(aid))
        return out

    def align_tasks(self, blob, seg_off, seg_len, seg_ad):
        out = np.empty(len(seg_ad), dtype=self.dtype)
        for t in range(len(seg_ad)):
            out[t] = self._one(bytes(blob[seg_off[t]:seg_off[t] + seg_len[t]]), int(seg_ad[t]))
        return out


class ReducedOracleEngine(OracleEngine):
    """The reduced-output calls (pcb_align_windows_best / _hits) answered from the oracle's full records: checks the host
    logic of the phase functions on those outputs without a GPU."""

    def align_windows_best(self, buf, lens, ids):
        from custom_porechop_abi_b200.engine import BEST_DTYPE
        rec = self.align_windows(buf, lens, ids)
        out = np.zeros(len(ids), dtype=BEST_DTYPE)
        for k in range(len(ids)):
            best = None
            for i in range(len(lens)):
                r = rec[i, k]
                if r['read_start'] != -1 and (best is None or int(r['matches']) * int(best['full_len']) > int(best['matches']) * int(r['full_len'])):
                    best = r
            if best is not None:
                out[k] = (best['matches'], best['full_len'])
        return out

    def align_windows_hits(self, buf, lens, ids, end_threshold, at_end, end_size, min_trim_size, want_ident=None, hits_cap=None,
                           hits_buf=None, ident_buf=None):
        import ctypes
        from custom_porechop_abi_b200.engine import HIT_DTYPE, WindowHits, threshold_e6
        rec = self.align_windows(buf, lens, ids)
        t6 = threshold_e6(end_threshold)
        hits = []
        for i in range(len(lens)):
            for k in range(len(ids)):
                r = rec[i, k]
                failed = r['read_start'] == -1
                if not failed and r['aligned_len'] <= 0:
                    continue
                num, den = (0, 1) if failed else (200000000 * int(r['matches']), int(r['aligned_len']))
                rs, re = int(r['read_start']), (0 if failed else int(r['read_end']) + 1)
                if num >= (2 * t6 + 1) * den and (rs != 0 if at_end else re != end_size) and re - rs >= min_trim_size:
                    hits.append((i, k, r['read_start'], r['read_end'], r['raw_score'], r['matches'], r['aligned_len'], r['full_len']))
        cols = np.flatnonzero(want_ident) if want_ident is not None else np.zeros(0, dtype=np.int64)
        ident = np.zeros((len(lens), len(cols)), dtype=np.uint16)
        for j, k in enumerate(cols):
            ok = rec['read_start'][:, k] != -1
            ident[:, j] = np.where(ok, (rec['matches'][:, k] << 9) | rec['full_len'][:, k], 0)
        return WindowHits(np.array(hits, dtype=HIT_DTYPE), ctypes.c_int64(len(hits)), ident)


def cases():
    with open(GOLDEN) as f:
        return json.load(f)['cases']


_OUT = {}


def out_golden():
    if not _OUT:
        with open(OUT_GOLDEN) as f:
            _OUT.update(json.load(f))
    return _OUT


def check_output(idx, reads, tmp_path):
    """What the reference writes for reads in this state (tests/golden/output_golden.json: its own get_fasta / get_fastq per
    read under 16 option sets, its filter_reads_by_adapter, and four output_reads runs -- one file, one FASTA file, barcode
    bins, untrimmed bins -- as hashes) against read_output.py."""
    import hashlib
    from custom_porechop_abi_b200 import read_output as ro
    g = out_golden()['cases'][idx]
    sha = lambda text: hashlib.sha1(text.encode()).hexdigest() if text else ''
    for v in g['records']:
        fn = ro.get_fasta if v['format'] == 'fasta' else ro.get_fastq
        texts = [fn(r, v['min_split'], v['discard_middle'], v['untrimmed']) for r in reads]
        bad = [r.name for r, t, want in zip(reads, texts, v['sha1']) if sha(t) != want]
        assert not bad and sum(len(t) for t in texts) == v['n_chars'], (v['format'], v['min_split'], v['discard_middle'], v['untrimmed'], bad)
    kept = ro.filter_reads_by_adapter(reads)
    assert [reads.index(r) for r in kept] == g['filtered']
    for tag, f in g['files'].items():
        a = f['args']
        d = tmp_path / ('out_%d_%s' % (idx, tag))
        d.mkdir()
        output = str(d / a['output']) if a['output'] else None
        bdir = str(d / a['barcode_dir']) if a['barcode_dir'] else None
        ro.output_reads(kept, a['out_format'], output, g['read_type'], 0, a['discard_middle'], a['min_split'], None, bdir,
                        g['fastq'], a['untrimmed'], 1, a['discard_unassigned'])
        where = bdir if bdir else str(d)
        got = {fn: hashlib.sha256(open(os.path.join(where, fn), 'rb').read()).hexdigest() for fn in sorted(os.listdir(where))}
        assert got == f['sha256'], (tag, sorted(got), sorted(f['sha256']))


def run_case(c, eng, read_blob=None):
    from custom_porechop_abi_b200 import porechop_batch as pb
    p = c['params']
    quals = out_golden()['quals'][c['fastq']]
    reads = [Read(n, s, q) for (n, s), q in zip(c['reads'], quals)]
    for r, alb in zip(reads, c.get('albacore', [None] * len(reads))):
        r.albacore_barcode_call = alb                       # Albacore's bin, from the directory the read came from
    search = [AdapterSet(d) for d in c['search_adapters']]
    # phase A
    matching = pb.find_matching_adapter_sets(reads, 0, p['end_size'], p['scoring'], None, p['adapter_threshold'], 1, search, engine=eng)
    assert [[a.best_start_score, a.best_end_score] for a in search] == c['phase_a']['best']
    assert [a.name for a in matching] == c['phase_a']['matching']
    # phases B and C on the sets the reference went on with (its 1D^2 / full-barcode rewrites are host logic)
    used = [AdapterSet(d) for d in c['matching_used']]
    pb.find_adapters_at_read_ends(reads, used, 0, p['end_size'], p['extra_end_trim'], p['end_threshold'], p['scoring'], None,
                                  p['min_trim_size'], 1, c['check_barcodes'], p['barcode_threshold'], p['barcode_diff'],
                                  c['require_two_barcodes'], c['forward_or_reverse'], engine=eng, read_blob=read_blob)
    for r, g in zip(reads, c['phase_b']):
        assert r.start_trim_amount == g['start_trim_amount'] and r.end_trim_amount == g['end_trim_amount'], r.name
        assert [[a.name, f, pa, s, e] for a, f, pa, s, e in r.start_adapter_alignments] == g['start_alignments'], r.name
        assert [[a.name, f, pa, s, e] for a, f, pa, s, e in r.end_adapter_alignments] == g['end_alignments'], r.name
        assert r.start_barcode_scores == g['start_barcode_scores'] and r.end_barcode_scores == g['end_barcode_scores'], r.name
        assert list(r.start_barcode_scores) == list(g['start_barcode_scores']) and list(r.end_barcode_scores) == list(g['end_barcode_scores']), r.name
        if c['check_barcodes']:
            assert r.barcode_call == g['barcode_call'], (r.name, r.barcode_call, g['barcode_call'])
    pb.find_adapters_in_read_middles(reads, used, 0, p['middle_threshold'], p['good_side'], p['bad_side'], p['scoring'], None, 1,
                                     False, engine=eng, read_blob=read_blob)
    for r, g in zip(reads, c['phase_c']):
        assert sorted(r.middle_adapter_positions) == g['middle_adapter_positions'], r.name
        assert sorted(r.middle_trim_positions) == g['middle_trim_positions'], r.name
        assert r.middle_hit_str == g['middle_hit_str'], r.name
        # the fork's read filter (porechop_abi.py:36-39, nanopore_read.py:23-30): an adapter alignment at both ends
        assert (bool(r.start_adapter_alignments) and bool(r.end_adapter_alignments)) == g['adapters_found'], r.name
    return reads


def test_golden_has_hits():
    cs = cases()
    assert len(cs) >= 13
    assert sum(r['barcode_call'] != 'none' for c in cs if c['check_barcodes'] for r in c['phase_b']) >= 20
    assert any(a for c in cs for a in c.get('albacore', []))
    assert {True, False} <= {r['adapters_found'] for c in cs for r in c['phase_c']}
    assert sum(len(r['middle_adapter_positions']) > 0 for c in cs for r in c['phase_c']) >= 10
    assert sum(r['start_trim_amount'] > 0 for c in cs for r in c['phase_b']) >= 10
    assert any(r['start_barcode_scores'] for c in cs for r in c['phase_b'])


@pytest.mark.parametrize('idx', range(13))
def test_phase_logic_with_oracle_engine(idx, tmp_path):
    reads = run_case(cases()[idx], OracleEngine())
    check_output(idx, reads, tmp_path)             # ... and what is written for reads in that state


@pytest.mark.parametrize('idx', range(13))
def test_phase_logic_on_reduced_outputs(idx):
    run_case(cases()[idx], ReducedOracleEngine())


def test_trimmed_bounds_follow_python_slicing():
    """Phase C addresses the trimmed read as a segment of the untrimmed one: same bases as the reference's slice, also
    when a trim exceeds the read (its stop index goes negative and Python counts it from the end)."""
    import random
    from custom_porechop_abi_b200.porechop_batch import _trimmed_bounds
    rng = random.Random(4)
    for _ in range(3000):
        r = Read('x', ''.join(rng.choice('ACGT') for _ in range(rng.randrange(0, 30))))
        r.start_trim_amount, r.end_trim_amount = rng.randrange(0, 40), rng.randrange(0, 40)
        lo, hi = _trimmed_bounds(r)
        assert r.seq[lo:hi] == r.get_seq_with_start_end_adapters_trimmed(), (r.seq, r.start_trim_amount, r.end_trim_amount)


def _write_fastq(path, records, lower=False, crlf=False):
    eol = '\r\n' if crlf else '\n'
    with open(path, 'w', newline='') as f:
        for name, seq in records:
            f.write('@' + name + ' extra words' + eol + (seq.lower() if lower else seq) + eol + '+' + eol + 'I' * len(seq) + eol)


def test_read_blob_matches_the_reference_loader_semantics(tmp_path):
    """ReadBlob (read_blob.py): the records of a FASTQ / FASTA file as one buffer + segments.  Names and sequences as the
    reference's loader + NanoporeRead give them (misc.py:123-168, nanopore_read.py:36-39), windows as _windows packs them,
    trimmed segments as Python slices them."""
    import gzip
    import random
    from custom_porechop_abi_b200 import porechop_batch as pb
    from custom_porechop_abi_b200.read_blob import load_read_blob
    records = cases()[0]['reads'] + [['empty', ''], ['rna', 'ACGUUUUACG'], ['n', 'ACGTNNNACGT']]
    reads = [Read(n, s) for n, s in records]
    for variant, kw in (('plain', {}), ('lower', {'lower': True}), ('crlf', {'crlf': True})):
        path = str(tmp_path / (variant + '.fastq'))
        _write_fastq(path, records, **kw)
        if variant == 'plain':                               # gzip goes through the same parser
            with open(path, 'rb') as f, gzip.open(path + '.gz', 'wb') as g:
                g.write(f.read())
            path = path + '.gz'
        blob = load_read_blob(path)
        assert len(blob) == len(reads) and blob.file_type == 'FASTQ'
        for i, r in enumerate(reads):
            assert blob.seq(i) == r.seq and blob.name(i) == r.name and blob.full_name(i) == r.name + ' extra words'
            assert blob.quals(i) == 'I' * len(r.seq)
        for at_end in (False, True):
            for end_size in (150, 100, 7):
                b1, l1 = blob.windows(end_size, at_end)
                b2, l2 = pb._windows(reads, end_size, at_end)
                assert (l1 == l2).all() and bytes(b1).upper().replace(b'U', b'T') == bytes(b2).replace(b'U', b'T')
    rng = random.Random(9)
    st = [rng.randrange(0, 300) if rng.random() < 0.7 else 0 for _ in reads]
    en = [rng.randrange(0, 300) if rng.random() < 0.7 else 0 for _ in reads]
    off, ln = blob.trimmed_segments(st, en)
    for i, r in enumerate(reads):
        r.start_trim_amount, r.end_trim_amount = st[i], en[i]
        got = blob.buf[off[i]:off[i] + ln[i]].tobytes().decode().upper()
        assert got.replace('U', 'T') == r.get_seq_with_start_end_adapters_trimmed().replace('U', 'T'), i
    # FASTA, sequences over several lines, blank lines in between
    path = str(tmp_path / 'x.fasta')
    with open(path, 'w') as f:
        for name, seq in records:
            f.write('>' + name + ' more\n')
            for k in range(0, len(seq), 60):
                f.write(seq[k:k + 60] + '\n')
            f.write('\n')
    blob = load_read_blob(path)
    assert len(blob) == len(reads) and blob.file_type == 'FASTA'
    for i, r in enumerate(reads):
        assert blob.seq(i) == r.seq and blob.name(i) == r.name


@pytest.mark.parametrize('idx', [0, 3, 5])
def test_phase_logic_with_read_blob(idx, tmp_path):
    """Phases B and C fed from a ReadBlob (windows gathered from the file buffer, phase C's blob = the file's bytes)."""
    from custom_porechop_abi_b200.read_blob import load_read_blob
    c = cases()[idx]
    path = str(tmp_path / 'reads.fastq')
    _write_fastq(path, c['reads'], lower=(idx == 3))
    run_case(c, OracleEngine(), read_blob=load_read_blob(path))


@pytest.mark.parametrize('idx', [0, 3, 5, 9])
def test_phase_logic_with_blob_reads(idx, tmp_path):
    """The reads as BlobRead objects of the ReadBlob itself: trims live in the blob's arrays, alignment lists are built on
    demand from its hit arrays (read_blob.apply_end_hits); same state as the reference's objects end up with."""
    from custom_porechop_abi_b200 import porechop_batch as pb
    from custom_porechop_abi_b200.read_blob import load_read_blob
    c = cases()[idx]
    p = c['params']
    path = str(tmp_path / 'reads.fastq')
    _write_fastq(path, c['reads'])
    blob = load_read_blob(path)
    reads = blob.make_reads()
    for r, alb in zip(reads, c.get('albacore', [None] * len(reads))):
        r.albacore_barcode_call = alb
    used = [AdapterSet(d) for d in c['matching_used']]
    eng = ReducedOracleEngine()
    pb.find_adapters_at_read_ends(reads, used, 0, p['end_size'], p['extra_end_trim'], p['end_threshold'], p['scoring'], None,
                                  p['min_trim_size'], 1, c['check_barcodes'], p['barcode_threshold'], p['barcode_diff'],
                                  c['require_two_barcodes'], c['forward_or_reverse'], engine=eng, read_blob=blob)
    for r, g in zip(reads, c['phase_b']):
        assert r.start_trim_amount == g['start_trim_amount'] and r.end_trim_amount == g['end_trim_amount'], r.name
        assert [[a.name, f, pa, s, e] for a, f, pa, s, e in r.start_adapter_alignments] == g['start_alignments'], r.name
        assert [[a.name, f, pa, s, e] for a, f, pa, s, e in r.end_adapter_alignments] == g['end_alignments'], r.name
        if c['check_barcodes']:
            assert r.barcode_call == g['barcode_call'], r.name
    pb.find_adapters_in_read_middles(reads, used, 0, p['middle_threshold'], p['good_side'], p['bad_side'], p['scoring'], None, 1,
                                     False, engine=eng, read_blob=blob)
    for r, g in zip(reads, c['phase_c']):
        assert sorted(r.middle_adapter_positions) == g['middle_adapter_positions'] and r.middle_hit_str == g['middle_hit_str'], r.name
        assert r.adapters_found() == g['adapters_found'], r.name


def _write_records(path, c):
    """The case's reads as the fixture holds them: full header, sequence, qualities (FASTA: 70 per line)."""
    quals = out_golden()['quals'][c['fastq']]
    fasta = c['fastq'].endswith('.fasta')
    with open(path, 'w') as f:
        for (name, seq), q in zip(c['reads'], quals):
            if fasta:
                f.write('>' + name + '\n' + ''.join(seq[i:i + 70] + '\n' for i in range(0, len(seq), 70)))
            else:
                f.write('@' + name + '\n' + seq + '\n+\n' + q + '\n')


def blob_reads_through_phases(c, eng, tmp_path):
    from custom_porechop_abi_b200 import porechop_batch as pb
    from custom_porechop_abi_b200.read_blob import load_read_blob
    p = c['params']
    path = str(tmp_path / ('records.' + ('fasta' if c['fastq'].endswith('.fasta') else 'fastq')))
    _write_records(path, c)
    blob = load_read_blob(path)
    reads = blob.make_reads()
    assert [r.name for r in reads] == [n for n, _ in c['reads']]
    for r, alb in zip(reads, c.get('albacore', [None] * len(reads))):
        r.albacore_barcode_call = alb
    used = [AdapterSet(d) for d in c['matching_used']]
    pb.find_adapters_at_read_ends(reads, used, 0, p['end_size'], p['extra_end_trim'], p['end_threshold'], p['scoring'], None,
                                  p['min_trim_size'], 1, c['check_barcodes'], p['barcode_threshold'], p['barcode_diff'],
                                  c['require_two_barcodes'], c['forward_or_reverse'], engine=eng, read_blob=blob)
    pb.find_adapters_in_read_middles(reads, used, 0, p['middle_threshold'], p['good_side'], p['bad_side'], p['scoring'], None, 1,
                                     False, engine=eng, read_blob=blob)
    return reads


def test_output_format_rules():
    """--format auto and the .gz rules of output_reads (porechop_abi.py:552-572), case by case as the reference decides them --
    including its quirk that an explicit 'fasta.gz' on stdout keeps the '.gz' and therefore takes the FASTQ branch (:643)."""
    from custom_porechop_abi_b200 import read_output as ro
    f = ro.resolve_format
    assert f('auto', None, 'FASTQ', None, 'in.fastq') == ('fastq', False)
    assert f('auto', None, 'FASTA', None, 'in.fasta.gz') == ('fasta', False)
    assert f('auto', None, 'FASTQ', 'bins', 'in.fastq.gz') == ('fastq', True)
    assert f('auto', None, 'FASTQ', 'bins', 'in.fastq') == ('fastq', False)
    assert f('auto', 'out.FASTA.gz', 'FASTQ', None, 'in.fastq') == ('fasta', True)
    assert f('auto', 'out.fastq.gz', 'FASTA', None, 'in.fasta') == ('fastq', True)
    assert f('auto', 'out.fasta', 'FASTQ', None, 'in.fastq') == ('fasta', False)
    assert f('auto', 'reads.fq', 'FASTQ', None, 'in.fastq') == ('fastq', False)
    assert f('auto', 'reads.txt', 'FASTA', None, 'in.fasta') == ('fasta', False)
    assert f('fasta.gz', 'x.out', 'FASTQ', None, 'in.fastq') == ('fasta', True)
    assert f('fasta.gz', None, 'FASTQ', None, 'in.fastq') == ('fasta.gz', False)
    r = Read('r', 'ACGT' * 5, 'I' * 20)
    recs = ro.records_of([r], 'fasta.gz', None, 'FASTQ', False, 1000, None, 'in.fastq', False, False)
    assert recs == [('', '@r\n' + 'ACGT' * 5 + '\n+\n' + 'I' * 20 + '\n', 0)]


def test_output_corner_cases(tmp_path):
    """Hand-made reads run through the reference's own get_fasta / get_fastq / get_split_read_parts (output_golden.json:
    'edge'): RNA, names with tab / space / neither, short quality strings, trims longer than the read, trim positions
    outside the read, min_split 0 -- on reference-shaped objects and on BlobRead objects of a file that holds them."""
    from custom_porechop_abi_b200 import read_output as ro
    from custom_porechop_abi_b200.read_blob import load_read_blob
    edge = out_golden()['edge']
    path = str(tmp_path / 'edge.fastq')
    with open(path, 'w') as f:
        for e in edge:
            f.write('@' + e['name'] + '\n' + e['seq'] + '\n+\n' + e['quals'] + '\n')
    blob_reads = load_read_blob(path).make_reads()
    for e, b in zip(edge, blob_reads):
        r = Read(e['name'], e['seq'], e['quals'])
        for x in (r, b):
            x.start_trim_amount, x.end_trim_amount = e['start_trim'], e['end_trim']
            x.middle_trim_positions = set(e['middle_trim_positions'])
            for min_split, parts in e['parts']:
                assert [list(p) for p in ro.get_split_read_parts(x, min_split)] == parts, (e['name'], min_split, type(x).__name__)
            for fmt, min_split, discard_middle, untrimmed, want in e['outputs']:
                got = (ro.get_fasta if fmt == 'fasta' else ro.get_fastq)(x, min_split, discard_middle, untrimmed)
                assert got == want, (e['name'], fmt, min_split, discard_middle, untrimmed, type(x).__name__)


@pytest.mark.parametrize('idx', range(13))
def test_output_of_blob_reads(idx, tmp_path):
    """File -> ReadBlob -> BlobRead objects -> phases B and C -> the records and files the reference writes, byte for byte."""
    from custom_porechop_abi_b200 import read_output as ro
    reads = blob_reads_through_phases(cases()[idx], ReducedOracleEngine(), tmp_path)
    check_output(idx, reads, tmp_path)
    for r in reads:                                # the methods the reference's own output_reads calls on a read
        assert r.get_fastq(20, False) == ro.get_fastq(r, 20, False) and r.get_fasta(1000, True, True) == ro.get_fasta(r, 1000, True, True)
        assert r.seq_length_with_start_end_adapters_trimmed() == len(r.get_seq_with_start_end_adapters_trimmed())
        assert r.get_split_read_parts(20) == ro.get_split_read_parts(r, 20)


def _write_input(c, tmp_path):
    """The case's reads as input of pipeline.trim: one file, or (Albacore cases) a directory tree whose sorted file order is
    the reads' order and whose directory names carry Albacore's bins (porechop_abi.py:190-197)."""
    alb = c.get('albacore', [None] * len(c['reads']))
    if not any(a is not None for a in alb):
        path = str(tmp_path / ('input.' + ('fasta' if c['fastq'].endswith('.fasta') else 'fastq')))
        _write_records(path, c)
        return path
    quals = out_golden()['quals'][c['fastq']]
    root = tmp_path / 'albacore'
    run, k = [], 0
    for i in range(len(alb) + 1):
        if i == len(alb) or (run and alb[i] != alb[run[0]]):
            call = alb[run[0]]
            sub = 'unclassified' if call == 'none' else ('barcode' + call[2:] if call else 'other')
            d = root / ('%03d' % k) / sub
            d.mkdir(parents=True)
            with open(str(d / 'reads.fastq'), 'w') as f:
                for j in run:
                    f.write('@' + c['reads'][j][0] + '\n' + c['reads'][j][1] + '\n+\n' + quals[j] + '\n')
            run, k = [], k + 1
        if i < len(alb):
            run.append(i)
    return str(root)


def check_pipeline(idx, eng, tmp_path):
    """pipeline.trim from the input file / directory to the output file and the barcode bins, against the bytes the reference's
    own main() steps wrote (output_golden.json: 'files'); the sets and the barcode direction the reference went on with after
    phase A come from the golden (its rewrites depend on its adapter database)."""
    import hashlib
    from custom_porechop_abi_b200 import pipeline
    c, g = cases()[idx], out_golden()['cases'][idx]
    p = c['params']
    src = _write_input(c, tmp_path)

    def between(matching, check_barcodes):
        assert [a.name for a in matching] == c['phase_a']['matching']
        return [AdapterSet(d) for d in c['matching_used']], c['forward_or_reverse']

    for tag in ('file_auto', 'file_fasta') + (('bins',) if c['check_barcodes'] else ()):
        a = g['files'][tag]['args']
        d = tmp_path / ('pipe_' + tag)
        d.mkdir()
        res = pipeline.trim(src, [AdapterSet(x) for x in c['search_adapters']], output=str(d / a['output']) if a['output'] else None,
                            out_format=a['out_format'], barcode_dir=str(d / a['barcode_dir']) if a['barcode_dir'] else None,
                            between_phases=between, engine=eng, scoring_scheme_vals=p['scoring'], end_size=p['end_size'],
                            adapter_threshold=p['adapter_threshold'], end_threshold=p['end_threshold'], extra_end_trim=p['extra_end_trim'],
                            min_trim_size=p['min_trim_size'], middle_threshold=p['middle_threshold'],
                            extra_middle_trim_good_side=p['good_side'], extra_middle_trim_bad_side=p['bad_side'],
                            discard_middle=a['discard_middle'], min_split_read_size=a['min_split'], barcode_threshold=p['barcode_threshold'],
                            barcode_diff=p['barcode_diff'], require_two_barcodes=c['require_two_barcodes'],
                            discard_unassigned=a['discard_unassigned'], untrimmed=a['untrimmed'])
        where = str(d / a['barcode_dir']) if a['barcode_dir'] else str(d)
        got = {fn: hashlib.sha256(open(os.path.join(where, fn), 'rb').read()).hexdigest() for fn in sorted(os.listdir(where))}
        assert got == g['files'][tag]['sha256'], (tag, sorted(got), sorted(g['files'][tag]['sha256']))
        assert [r.albacore_barcode_call for r in res['reads']] == c.get('albacore', [None] * len(c['reads']))


@pytest.mark.parametrize('idx', [0, 2, 3, 5, 8, 9, 12])
def test_pipeline_from_file_to_file(idx, tmp_path):
    check_pipeline(idx, ReducedOracleEngine(), tmp_path)


@pytest.mark.gpu
def test_phases_on_gpu(engine, tmp_path):
    from custom_porechop_abi_b200.read_blob import load_read_blob
    for k, c in enumerate(cases()):
        da, db = tmp_path / ('a%d' % k), tmp_path / ('b%d' % k)
        da.mkdir(); db.mkdir()
        check_output(k, run_case(c, engine), da)                              # reference-shaped objects -> records and files
        check_output(k, blob_reads_through_phases(c, engine, db), db)         # file -> ReadBlob -> BlobRead -> the same bytes
    for k in (0, 5, 9, 12):                                                   # ... and the whole way from input path to output files
        d = tmp_path / ('p%d' % k)
        d.mkdir()
        check_pipeline(k, engine, d)
        path = str(tmp_path / ('reads%d.fastq' % k))
        _write_fastq(path, c['reads'], lower=(k % 2 == 1))
        run_case(c, engine, read_blob=load_read_blob(path))
