"""check_compatibility (porechop_abi/ab_initio_src/compatibility.cpp:160-168), the other call site of the alignment in the
reference: golden flags made with the reference's own library (tests/golden/make_compat_golden.py), the oracle's
restatement against them (CPU), and the CUDA library against both (GPU, through the C ABI)."""
import json
import os
import random

import numpy as np
import pytest

from oracle_lib import COMPAT_REF_SO, mutate, oracle_compat, reference_compat

GOLDEN = os.path.join(os.path.dirname(__file__), 'golden', 'compat_golden.json')


def cases():
    with open(GOLDEN) as f:
        return json.load(f)['cases']


def test_golden_covers_all_flags():
    cs = cases()
    assert len(cs) >= 500 and all(sum(1 for c in cs if c[2] == k) >= 40 for k in range(3))


def test_oracle_restatement_matches_the_golden_flags():
    bad = [(a, b, f, oracle_compat(a, b)) for a, b, f in cases() if oracle_compat(a, b) != f]
    assert not bad, bad[:5]


@pytest.mark.skipif(not os.path.isfile(COMPAT_REF_SO), reason='reference library not built (no /root/reference here)')
def test_oracle_restatement_matches_the_reference_library():
    rng = random.Random(5)
    for _ in range(1500):
        a = ''.join(rng.choice('ACGT') for _ in range(rng.choice([5, 12, 24, 31, 50])))
        b = mutate(rng, a, rng.choice([0, 0.1, 0.3])) if rng.random() < 0.7 else ''.join(rng.choice('ACGT') for _ in range(rng.randrange(1, 40)))
        b = (''.join(rng.choice('ACGT') for _ in range(rng.randrange(0, 9))) + b)[:60] or 'C'
        assert oracle_compat(a, b) == reference_compat(a, b), (a, b)


@pytest.mark.gpu
def test_check_compatibility_on_gpu(engine):
    from custom_porechop_abi_b200.cpp_function_wrappers import check_compatibility
    cs = cases()
    seqs = [c[0] for c in cs] + [c[1] for c in cs]
    pairs = np.array([[i, len(cs) + i] for i in range(len(cs))], dtype=np.int32)
    flags = engine.check_compatibility(seqs, pairs)
    bad = [(cs[i][0], cs[i][1], cs[i][2], int(flags[i])) for i in range(len(cs)) if int(flags[i]) != cs[i][2]]
    assert not bad, bad[:5]
    for a, b, f in cs[:40]:                                   # the drop-in symbol, one pair per call
        assert check_compatibility(a, b) == f, (a, b)
    # all-vs-all of a family of candidates, as consensus.py builds its matrix; against the oracle
    rng = random.Random(8)
    base = ''.join(rng.choice('ACGT') for _ in range(30))
    fam = [base] + [mutate(rng, base, 0.08)[rng.randrange(0, 4):] for _ in range(12)] + [''.join(rng.choice('ACGT') for _ in range(25)) for _ in range(4)]
    fam = [s or 'A' for s in fam]
    allp = np.array([[i, j] for i in range(len(fam)) for j in range(len(fam))], dtype=np.int32)
    got = engine.check_compatibility(fam, allp)
    for (i, j), g in zip(allp.tolist(), got.tolist()):
        assert g == oracle_compat(fam[i], fam[j]), (fam[i], fam[j])
    engine.set_scoring((3, -6, -5, -2))
