"""CPU: pins the oracle. (1) against the golden vectors generated with the reference's own library,
(2) against that library directly when oracle/_ref is present, on seeded random and adversarial inputs."""
import json
import os
import random

import pytest

from oracle_lib import oracle_align, oracle_string, random_case, reference_available, reference_string

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden', 'adapter_alignment_golden.json')
SCHEMES = [(3, -6, -5, -2)] * 3 + [(3, -6, -5, -5), (1, -1, -1, -1), (2, -3, -4, -1), (5, -4, -10, -1), (1, -1, -3, -2),
                                   (3, -6, -2, -5), (1, 0, 0, 0), (1, -1, 1, 1), (0, 0, 0, 0), (2, 2, -1, -2), (3, -6, 1, -2)]


def golden_cases():
    with open(GOLDEN) as f:
        return json.load(f)['cases']


def test_oracle_matches_golden_vectors():
    cases = golden_cases()
    assert len(cases) > 500
    for c in cases:
        assert oracle_string(c['read'], c['adapter'], c['scores']) == c['expect'], (c['read'][:60], c['adapter'])


def test_empty_inputs_are_the_failed_record():
    for r, a in (('', 'ACGT'), ('ACGT', ''), ('', '')):
        o = oracle_align(r, a)
        assert o.read_start == -1
        assert oracle_string(r, a) == '-1,0,-1,0,-2147483648,0.000000,0.000000'


def test_match_counts_are_equal():
    # alignedMatchCount == fullAdapterMatchCount for every alignment, which is why the
    # device record carries one 'matches' field
    rng = random.Random(3)
    for _ in range(4000):
        r, a = random_case(rng)
        o = oracle_align(r, a, rng.choice(SCHEMES))
        assert o.aligned_matches == o.full_matches


@pytest.mark.skipif(not reference_available(), reason='oracle/_ref not built (needs /root/reference)')
@pytest.mark.parametrize('seed', [1, 2, 3, 4])
def test_oracle_matches_reference_library(seed):
    rng = random.Random(seed)
    for _ in range(6000):
        r, a = random_case(rng)
        s = rng.choice(SCHEMES)
        assert oracle_string(r, a, s) == reference_string(r, a, s), (r, a, s)


@pytest.mark.skipif(not reference_available(), reason='oracle/_ref not built (needs /root/reference)')
def test_oracle_matches_reference_on_long_reads():
    rng = random.Random(77)
    ad = 'AATGTACTTCGTTCAGTTACGTATTGCT'
    for m in (1000, 5000, 20000):
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        p = rng.randrange(0, m - 40)
        r = r[:p] + ad[:20] + 'T' + ad[20:] + r[p + 29:]
        assert oracle_string(r, ad) == reference_string(r, ad)
