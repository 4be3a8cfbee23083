#!/usr/bin/env python3
"""
Generates tests/golden/phases_golden.json with the REFERENCE ITSELF: the reference's own Python
(porechop_abi/porechop_abi.py: find_matching_adapter_sets, find_adapters_at_read_ends,
find_adapters_in_read_middles, driving nanopore_read.NanoporeRead) on top of the reference's own C++ library
(oracle/_ref/cpp_functions_ref.so, built by oracle/Makefile), run single-threaded on the reference's own test reads.
The package is imported from a scratch copy (the reference tree is read-only and has no built library); nothing of
it is copied into the repo -- only inputs (reads, adapter sets) and the resulting per-read / per-set state, as JSON.
Run in the build container only:  python tests/golden/make_phase_golden.py
"""
import hashlib
import json
import os
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = '/root/reference'
REF_SO = os.path.join(ROOT, 'oracle', '_ref', 'cpp_functions_ref.so')

DEFAULTS = dict(end_size=150, scoring=[3, -6, -5, -2], adapter_threshold=90.0, end_threshold=75.0, extra_end_trim=2,
                min_trim_size=4, middle_threshold=90.0, good_side=10, bad_side=100, barcode_threshold=75.0,
                barcode_diff=5.0)


def load_records(path, pc):
    """(name, sequence, quals, albacore call) of every read of a file or Albacore directory, through the reference's own
    loader (porechop_abi.py:133-187)."""
    reads, _, _ = pc.load_reads(path, 0, open(os.devnull, 'w'), 10000)
    return [(r.name, r.seq, r.quals, r.albacore_barcode_call) for r in reads]


def adapter_dict(a):
    return {'name': a.name, 'start_sequence': list(a.start_sequence) if a.start_sequence else None,
            'end_sequence': list(a.end_sequence) if a.end_sequence else None}


def main():
    tmp = tempfile.mkdtemp(prefix='pc_ref_')
    shutil.copytree(os.path.join(REF, 'porechop_abi'), os.path.join(tmp, 'porechop_abi'),
                    ignore=shutil.ignore_patterns('include', 'src', '*.so', '__pycache__'))
    shutil.copy(REF_SO, os.path.join(tmp, 'porechop_abi', 'cpp_functions.so'))
    sys.path.insert(0, tmp)
    import warnings
    warnings.simplefilter('ignore')
    from porechop_abi import porechop_abi as pc
    from porechop_abi import adapters as ad
    from porechop_abi.nanopore_read import NanoporeRead

    devnull = open(os.devnull, 'w')
    P = DEFAULTS
    cases = []
    out_cases, quals_of = [], {}
    sha = lambda text: hashlib.sha1(text.encode()).hexdigest() if text else ''
    for fastq, check_barcodes, variants in (
            ('test_one_adapter_set.fastq', False, [dict(), dict(min_trim_size=7, end_size=100), dict(middle_threshold=80.0, extra_end_trim=0)]),
            ('test_two_adapter_sets.fastq', False, [dict()]),
            ('test_barcodes.fastq', True, [dict(), dict(require_two=True)]),
            ('test_choose_barcodes_1.fasta', True, [dict()]),
            ('test_choose_barcodes_2.fasta', True, [dict(), dict(require_two=True, barcode_diff=1.0)]),
            ('test_albacore_directory', True, [dict(), dict(barcode_threshold=60.0)]),
            ('test_format.fastq', False, [dict()]),
            ('test_format_barcodes.fasta', True, [dict(end_threshold=70.5)])):
        for var in variants:
            p = dict(P); p.update({k: v for k, v in var.items() if k != 'require_two'})
            require_two = var.get('require_two', False)
            records = load_records(os.path.join(REF, 'test', fastq), pc)
            reads = [NanoporeRead(n, s, q) for n, s, q, _ in records]
            for r, rec in zip(reads, records):
                r.albacore_barcode_call = rec[3]
            for a in ad.ADAPTERS:                          # the reference keeps scores on module-level objects
                a.best_start_score, a.best_end_score = 0.0, 0.0
            search = [a for a in ad.ADAPTERS if '(full sequence)' not in a.name]
            matching = pc.find_matching_adapter_sets(reads, 0, p['end_size'], p['scoring'], devnull, p['adapter_threshold'], 1)
            phase_a = {'best': [[a.best_start_score, a.best_end_score] for a in search],
                       'matching': [a.name for a in matching]}
            # the reference's own order between the phases (porechop_abi.py:78-88)
            matching = pc.fix_up_1d2_sets(matching)
            forward_or_reverse = pc.choose_barcoding_kit(matching, 0, devnull) if check_barcodes else None
            matching = pc.add_full_barcode_adapter_sets(matching)
            pc.find_adapters_at_read_ends(reads, matching, 0, p['end_size'], p['extra_end_trim'], p['end_threshold'],
                                          p['scoring'], devnull, p['min_trim_size'], 1, check_barcodes,
                                          p['barcode_threshold'], p['barcode_diff'], require_two, forward_or_reverse)
            phase_b = [{'start_trim_amount': r.start_trim_amount, 'end_trim_amount': r.end_trim_amount,
                        'start_alignments': [[a.name, f, pa, s, e] for a, f, pa, s, e in r.start_adapter_alignments],
                        'end_alignments': [[a.name, f, pa, s, e] for a, f, pa, s, e in r.end_adapter_alignments],
                        'start_barcode_scores': r.start_barcode_scores, 'end_barcode_scores': r.end_barcode_scores,
                        'barcode_call': r.barcode_call} for r in reads]
            # the fork's read filter (porechop_abi.py:36-39 -> nanopore_read.py:23-30) looks at the state phase B leaves
            pc.find_adapters_in_read_middles(reads, matching, 0, p['middle_threshold'], p['good_side'], p['bad_side'],
                                             p['scoring'], devnull, 1, False)
            phase_c = [{'middle_adapter_positions': sorted(r.middle_adapter_positions),
                        'middle_trim_positions': sorted(r.middle_trim_positions),
                        'middle_hit_str': r.middle_hit_str, 'adapters_found': bool(r.adapters_found())} for r in reads]
            # ---- the far side of the path: what the reference writes for these reads (output_golden.json) ----
            quals_of[fastq] = [q for _, _, q, _ in records]
            per_variant = []
            for fmt in ('fasta', 'fastq'):
                for min_split in (1000, 20):
                    for discard_middle in (False, True):
                        for untrimmed in (False, True):
                            texts = [(r.get_fasta if fmt == 'fasta' else r.get_fastq)(min_split, discard_middle, untrimmed) for r in reads]
                            per_variant.append({'format': fmt, 'min_split': min_split, 'discard_middle': discard_middle, 'untrimmed': untrimmed,
                                                'sha1': [sha(t) for t in texts], 'n_chars': sum(len(t) for t in texts)})
            filtered = pc.filter_reads_by_adapter(reads)
            files = {}
            out_dir = tempfile.mkdtemp(prefix='pc_out_')
            read_type = 'FASTA' if fastq.endswith('.fasta') else 'FASTQ'
            for tag, kw in (('file_auto', dict(out_format='auto', output=os.path.join(out_dir, 'out.fastq'), barcode_dir=None, untrimmed=False, discard_unassigned=False, discard_middle=False, min_split=20)),
                            ('file_fasta', dict(out_format='fasta', output=os.path.join(out_dir, 'out2.txt'), barcode_dir=None, untrimmed=False, discard_unassigned=False, discard_middle=True, min_split=1000)),
                            ('bins', dict(out_format='auto', output=None, barcode_dir=os.path.join(out_dir, 'bins'), untrimmed=False, discard_unassigned=False, discard_middle=False, min_split=20)),
                            ('bins_untrimmed', dict(out_format='fastq', output=None, barcode_dir=os.path.join(out_dir, 'bins2'), untrimmed=True, discard_unassigned=True, discard_middle=False, min_split=1000))):
                pc.output_reads(filtered, kw['out_format'], kw['output'], read_type, 0, kw['discard_middle'], kw['min_split'], devnull,
                                kw['barcode_dir'], os.path.join(REF, 'test', fastq), kw['untrimmed'], 1, kw['discard_unassigned'])
                got = {}
                if kw['barcode_dir']:
                    for fn in sorted(os.listdir(kw['barcode_dir'])):
                        got[fn] = hashlib.sha256(open(os.path.join(kw['barcode_dir'], fn), 'rb').read()).hexdigest()
                else:
                    got[os.path.basename(kw['output'])] = hashlib.sha256(open(kw['output'], 'rb').read()).hexdigest()
                files[tag] = {'args': {k: (os.path.basename(v) if isinstance(v, str) and os.sep in v else v) for k, v in kw.items()}, 'sha256': got}
            shutil.rmtree(out_dir)
            out_cases.append({'fastq': fastq, 'read_type': read_type, 'filtered': [reads.index(r) for r in filtered], 'records': per_variant, 'files': files})
            cases.append({'fastq': fastq, 'params': p, 'check_barcodes': check_barcodes, 'require_two_barcodes': require_two,
                          'forward_or_reverse': forward_or_reverse,
                          'reads': [[n, s] for n, s, _, _ in records], 'albacore': [a for _, _, _, a in records],
                          'search_adapters': [adapter_dict(a) for a in search],
                          'matching_used': [adapter_dict(a) for a in matching],
                          'phase_a': phase_a, 'phase_b': phase_b, 'phase_c': phase_c})
    out = os.path.join(HERE, 'phases_golden.json')
    with open(out, 'w') as f:
        json.dump({'generator': 'tests/golden/make_phase_golden.py', 'cases': cases}, f, separators=(',', ':'))
    # hand-made reads for the corners the fixtures do not reach: RNA, names with tab / space / neither, short quality strings,
    # trims longer than the read (Python's negative-index slicing decides), trim positions outside the read, min_split 0
    edge_in = [('r1 some desc', 'ACGU' * 30 + 'UUUU', 'I' * 124, 3, 5, []),
               ('r2\tx y', 'ACGT' * 50, '5' * 150, 0, 0, list(range(50, 61)) + [-3, 1000, 199]),
               ('r3', 'ACGT' * 10, '', 30, 30, []),
               ('r4 a b', 'ACGT' * 100, 'ABCDEFGH' * 50, 10, 395, [2, 3]),
               ('r5', 'acgtn' * 40, '!' * 200, 7, 0, [0, 1, 2, 3, 4, 5, 190, 191, 192, 100]),
               ('r6_u', 'AUGC' * 20, 'I' * 80, 0, 2, list(range(30, 40))),
               ('r7', 'ACGT' * 20, 'I' * 80, 0, 0, list(range(0, 80)))]
    edge = []
    for name, seq, quals, st, et, mid in edge_in:
        r = NanoporeRead(name, seq, quals)
        r.start_trim_amount, r.end_trim_amount, r.middle_trim_positions = st, et, set(mid)
        outs = []
        for fmt in ('fasta', 'fastq'):
            for min_split in (0, 20, 1000):
                for discard_middle in (False, True):
                    for untrimmed in (False, True):
                        outs.append([fmt, min_split, discard_middle, untrimmed,
                                     (r.get_fasta if fmt == 'fasta' else r.get_fastq)(min_split, discard_middle, untrimmed)])
        edge.append({'name': name, 'seq': seq, 'quals': quals, 'start_trim': st, 'end_trim': et, 'middle_trim_positions': mid,
                     'parts': [[min_split, r.get_split_read_parts(min_split)] for min_split in (0, 20, 1000)], 'outputs': outs})
    out2 = os.path.join(HERE, 'output_golden.json')
    with open(out2, 'w') as f:
        json.dump({'generator': 'tests/golden/make_phase_golden.py', 'quals': quals_of, 'cases': out_cases, 'edge': edge}, f, separators=(',', ':'))
    print('wrote %s: %d cases, %.0f KB' % (out2, len(out_cases), os.path.getsize(out2) / 1e3))
    shutil.rmtree(tmp)
    n_hits = sum(len(r['middle_adapter_positions']) > 0 for c in cases for r in c['phase_c'])
    print('wrote %s: %d cases, %d reads with middle hits, %.0f KB' % (out, len(cases), n_hits, os.path.getsize(out) / 1e3))


if __name__ == '__main__':
    main()
