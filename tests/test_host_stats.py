"""CPU: the statistics code the kernels run (csrc/align_stats.h, host build via pcb_stats_from_path) against the
oracle's ScoredAlignment restatement, on the oracle's own alignment paths."""
import ctypes
import random

from oracle_lib import oracle_align, oracle_path, random_case

SCHEMES = [(3, -6, -5, -2)] * 3 + [(3, -6, -5, -5), (1, -1, -1, -1), (2, -3, -4, -1), (1, 0, 0, 0), (1, -1, 1, 1),
                                   (0, 0, 0, 0), (2, 2, -1, -2), (3, -6, 1, -2)]


def test_stats_from_oracle_paths():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    rng = random.Random(11)
    for _ in range(6000):
        r, a = random_case(rng)
        s = rng.choice(SCHEMES)
        ops, flags, hh, hv, th, tv, score = oracle_path(r, a, s)
        res = w.PcbResult()
        rc = w.C_LIB.pcb_stats_from_path(ops, len(ops), hh, hv, th, tv, flags, score, ctypes.byref(res))
        assert rc == 0
        o = oracle_align(r, a, s)
        got = (res.read_start, res.read_end, res.adapter_start, res.adapter_end, res.raw_score, res.matches,
               res.matches, res.aligned_len, res.full_len)
        want = (o.read_start, o.read_end, o.adapter_start, o.adapter_end, o.raw_score, o.aligned_matches,
                o.full_matches, o.aligned_len, o.full_len)
        assert got == want, (r, a, s, ops, got, want)
