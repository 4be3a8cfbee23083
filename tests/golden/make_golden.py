#!/usr/bin/env python3
"""
Generates tests/golden/adapter_alignment_golden.json with the REFERENCE ITSELF: the reference's own
adapterAlignment (porechop_abi/src/adapter_align.cpp + src/alignment.cpp, compiled unmodified into
oracle/_ref/cpp_functions_ref.so by oracle/Makefile) is called on
  * end windows and whole reads of the reference's own test reads (test/test_one_adapter_set.fastq,
    test_two_adapter_sets.fastq, test_barcodes.fastq) against a sample of its adapter database
    (porechop_abi/adapters.py), exactly the pairs phases A/B/C of the reference align, and
  * hand-written edge cases (length-1 inputs, N / '-' symbols, lower case, U, overhangs at either end,
    empty inputs, linear and unusual scoring schemes).
Run in the build container only (needs /root/reference); the JSON it writes is what travels.
"""
import importlib.util
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import reference_string  # noqa: E402

REF = '/root/reference'


def load_fastq(path):
    seqs = []
    with open(path) as f:
        lines = f.read().split('\n')
    for i in range(0, len(lines) - 3, 4):
        seqs.append(lines[i + 1].strip().upper())
    return seqs


def main():
    spec = importlib.util.spec_from_file_location('adapters', os.path.join(REF, 'porechop_abi', 'adapters.py'))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rng = random.Random(2024)
    adapter_seqs = []
    for a in mod.ADAPTERS[:11] + rng.sample(mod.ADAPTERS[11:], 8):
        if a.start_sequence:
            adapter_seqs.append(a.start_sequence[1])
        if a.end_sequence:
            adapter_seqs.append(a.end_sequence[1])
    for maker in [getattr(mod, n) for n in dir(mod) if n.startswith("make_full_")]:
        try:
            fa = maker(3)
            adapter_seqs.append(fa.start_sequence[1])
            if fa.end_sequence:
                adapter_seqs.append(fa.end_sequence[1])
        except Exception:
            pass
    adapter_seqs = sorted(set(adapter_seqs))
    reads = []
    for fn in ('test_one_adapter_set.fastq', 'test_two_adapter_sets.fastq', 'test_barcodes.fastq'):
        reads += load_fastq(os.path.join(REF, 'test', fn))
    default = [3, -6, -5, -2]
    cases = []

    def add(read, adapter, scores):
        cases.append({'read': read, 'adapter': adapter, 'scores': list(scores), 'expect': reference_string(read, adapter, scores)})

    for r in reads:                       # phases A / B: 150-base end windows vs every sampled adapter
        for a in adapter_seqs:
            if rng.random() < 0.25:
                add(r[:150], a, default)
            if rng.random() < 0.25:
                add(r[-150:], a, default)
    for r in reads[:9]:                   # phase C: whole reads (2-12 kb) vs the adapters of the matching sets
        for a in adapter_seqs[:4]:
            add(r, a, default)
    # a masked read as find_middle_adapters produces it (nanopore_read.py:233-234)
    r = reads[5]
    add(r[:2000] + '-' * 30 + r[2030:], adapter_seqs[0], default)
    for r_, a_ in [("ACGTACGTTTGACCA", "GGGGGGG"), ("ACGTACGTTTGACCA", "TTGACCATTAGC"), ("ACGTACGTTTGACCA", "CCCACGTACG"),
                   ("ACGTACGTTTGACCA", "ACGTTTG"), ("A", "C"), ("A", "A"), ("ACGT", "ACGTACGTTTGACCA"), ("NNNN", "ACGT"),
                   ("acgtuu--xx", "ACGTTTNNNN"), ("", "ACGT"), ("ACGT", ""), ("", ""), ("ACGTNNACGT", "ACGTNNACGT"),
                   ("TTTTTTTTTT", "TTTT"), ("ACGTACGTAC", "ACGGTACG"), ("ACGTACGTAC", "ACTACG")]:
        for sc in (default, [3, -6, -5, -5], [1, -1, -3, -2], [2, -3, -4, -1], [1, 0, 0, 0], [1, -1, 1, 1], [3, -6, -2, -5]):
            add(r_, a_, sc)
    out = os.path.join(HERE, 'adapter_alignment_golden.json')
    with open(out, 'w') as f:
        json.dump({'generator': 'tests/golden/make_golden.py', 'reference_library': 'oracle/_ref/cpp_functions_ref.so '
                   '(g++ -O3 on porechop_abi/src/adapter_align.cpp + alignment.cpp, SeqAn 2.4.0 headers)',
                   'cases': cases}, f, separators=(',', ':'))
    print(len(cases), 'cases ->', out, os.path.getsize(out), 'bytes')


if __name__ == '__main__':
    main()
