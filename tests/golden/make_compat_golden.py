#!/usr/bin/env python3
"""
Generates tests/golden/compat_golden.json with the REFERENCE ITSELF: check_compatibility of the reference's
ab_initio_src/compatibility.cpp, compiled by oracle/Makefile into oracle/_ref/compatibility_ref.so, on pairs of candidate
adapters of the kind the ab-initio consensus step compares (consensus.py:72-100): a sequence against mutated, truncated,
shifted and embedded copies of itself, and against unrelated ones.  Run in the build container only.
"""
import json
import os
import random
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import mutate, reference_compat  # noqa: E402


def main():
    rng = random.Random(20260)
    rnd = lambda n: ''.join(rng.choice('ACGT') for _ in range(n))
    cases = []
    for _ in range(700):
        a = rnd(rng.choice([8, 16, 24, 28, 33, 40, 64, 90]))
        kind = rng.randrange(7)
        if kind == 0:
            b = mutate(rng, a, rng.choice([0.0, 0.05, 0.1, 0.2]))
        elif kind == 1:
            k = rng.randrange(0, len(a) // 2 + 1)
            b = mutate(rng, a[k:], rng.choice([0.0, 0.08])) + rnd(rng.randrange(0, 8))
        elif kind == 2:
            b = rnd(rng.randrange(0, 12)) + mutate(rng, a, rng.choice([0.0, 0.1])) + rnd(rng.randrange(0, 12))
        elif kind == 3:
            lo = rng.randrange(0, len(a) // 2); hi = rng.randrange(lo + 2, len(a) + 1)
            b = mutate(rng, a[lo:hi], rng.choice([0.0, 0.1]))
        elif kind == 4:
            b = rnd(rng.choice([8, 24, 40]))
        elif kind == 5:
            b = a[len(a) // 2:] + rnd(len(a) // 2)
        else:
            b = mutate(rng, a, 0.12).lower().replace('t', rng.choice(['t', 'u', 'n']))
        if not b:
            b = 'A'
        if rng.random() < 0.5:
            a, b = b, a
        cases.append([a, b, reference_compat(a, b)])
    out = os.path.join(HERE, 'compat_golden.json')
    with open(out, 'w') as f:
        json.dump({'generator': 'tests/golden/make_compat_golden.py', 'cases': cases}, f, separators=(',', ':'))
    hist = [sum(1 for c in cases if c[2] == k) for k in range(3)]
    print('wrote %s: %d pairs, flags 0/1/2 = %s' % (out, len(cases), hist))


if __name__ == '__main__':
    main()
