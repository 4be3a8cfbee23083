"""Test-side loader of the CPU oracle (oracle/), and of the compiled reference (oracle/_ref) when present."""
import ctypes
import os
import random
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, 'oracle')
ORACLE_SO = os.path.join(ORACLE_DIR, 'libporechop_oracle.so')
REF_SO = os.path.join(ORACLE_DIR, '_ref', 'cpp_functions_ref.so')

FIELDS = ('read_start', 'read_end', 'adapter_start', 'adapter_end', 'raw_score',
          'aligned_matches', 'aligned_len', 'full_matches', 'full_len')


class OracleResult(ctypes.Structure):
    _fields_ = [(f, ctypes.c_int) for f in FIELDS]


def _build():
    src = os.path.join(ORACLE_DIR, 'porechop_oracle.c')
    if (not os.path.isfile(ORACLE_SO)) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(src):
        subprocess.check_call(['gcc', '-O2', '-fPIC', '-shared', '-std=c11', '-o', ORACLE_SO, src])


_lib = None


def oracle():
    global _lib
    if _lib is None:
        _build()
        _lib = ctypes.CDLL(ORACLE_SO)
        _lib.pc_oracle_align_n.argtypes = [ctypes.c_char_p, ctypes.c_long, ctypes.c_char_p, ctypes.c_long] + \
            [ctypes.c_int] * 4 + [ctypes.POINTER(OracleResult)]
        _lib.pc_oracle_adapter_alignment.argtypes = [ctypes.c_char_p, ctypes.c_char_p] + [ctypes.c_int] * 4
        _lib.pc_oracle_adapter_alignment.restype = ctypes.c_void_p
        _lib.pc_oracle_free.argtypes = [ctypes.c_void_p]
        _lib.pc_oracle_path.argtypes = [ctypes.c_char_p, ctypes.c_long, ctypes.c_char_p, ctypes.c_long] + \
            [ctypes.c_int] * 4 + [ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(ctypes.c_long)]
        _lib.pc_oracle_path.restype = ctypes.c_long
    return _lib


def _b(s):
    return s.encode() if isinstance(s, str) else bytes(s)


def oracle_align(read, adapter, scores=(3, -6, -5, -2)):
    r = OracleResult()
    rb, ab = _b(read), _b(adapter)
    rc = oracle().pc_oracle_align_n(rb, len(rb), ab, len(ab), *scores, ctypes.byref(r))
    assert rc == 0
    return r


def oracle_string(read, adapter, scores=(3, -6, -5, -2)):
    lib = oracle()
    p = lib.pc_oracle_adapter_alignment(_b(read), _b(adapter), *scores)
    s = ctypes.cast(p, ctypes.c_char_p).value.decode()
    lib.pc_oracle_free(p)
    return s


def oracle_path(read, adapter, scores=(3, -6, -5, -2)):
    """(ops front-to-back, match flags, head_h, head_v, tail_h, tail_v, score)"""
    rb, ab = _b(read), _b(adapter)
    ops = ctypes.create_string_buffer(len(rb) + len(ab) + 2)
    flags = ctypes.create_string_buffer(len(rb) + len(ab) + 2)
    meta = (ctypes.c_long * 5)()
    n = oracle().pc_oracle_path(rb, len(rb), ab, len(ab), *scores, ops, flags, meta)
    assert n >= 0
    return ops.raw[:n], flags.raw[:n], meta[0], meta[1], meta[2], meta[3], meta[4]


_ref = None


def reference_available():
    return os.path.isfile(REF_SO)


def reference_string(read, adapter, scores=(3, -6, -5, -2)):
    """The reference's own adapterAlignment, compiled from /root/reference into oracle/_ref (oracle/Makefile)."""
    global _ref
    if _ref is None:
        _ref = ctypes.CDLL(REF_SO)
        _ref.adapterAlignment.argtypes = [ctypes.c_char_p, ctypes.c_char_p] + [ctypes.c_int] * 4
        _ref.adapterAlignment.restype = ctypes.c_void_p
        _ref.freeCString.argtypes = [ctypes.c_void_p]
    p = _ref.adapterAlignment(_b(read), _b(adapter), *scores)
    s = ctypes.cast(p, ctypes.c_char_p).value.decode()
    _ref.freeCString(p)
    return s


# ---- shared random-case generator ------------------------------------------------------------------
def mutate(rng, s, rate):
    out = []
    for ch in s:
        x = rng.random()
        if x < rate / 3:
            continue
        if x < 2 * rate / 3:
            out.append(rng.choice('ACGT')); out.append(ch); continue
        if x < rate:
            out.append(rng.choice('ACGT')); continue
        out.append(ch)
    return ''.join(out)


def random_case(rng, n_choices=(1, 2, 3, 5, 8, 12, 22, 24, 28, 33, 50), m_choices=(1, 2, 3, 7, 20, 60, 150, 150, 300),
                alphabets=('ACGT', 'ACGT', 'AC', 'ACGTN', 'A', 'ACGT-')):
    n = rng.choice(n_choices)
    m = rng.choice(m_choices)
    alpha = rng.choice(alphabets)
    a = ''.join(rng.choice(alpha) for _ in range(n))
    r = ''.join(rng.choice(alpha) for _ in range(m))
    if rng.random() < 0.6 and m > n:
        p = rng.randrange(-n, m)
        ins = mutate(rng, a, rng.choice([0, 0.05, 0.15, 0.3]))
        r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:max(m, 1)]
        if not r:
            r = 'A'
    return r, a


def rec_matches_oracle(rec, o):
    return (int(rec['read_start']) == o.read_start and int(rec['read_end']) == o.read_end and
            int(rec['adapter_start']) == o.adapter_start and int(rec['adapter_end']) == o.adapter_end and
            int(rec['raw_score']) == o.raw_score and int(rec['matches']) == o.aligned_matches and
            int(rec['matches']) == o.full_matches and int(rec['aligned_len']) == o.aligned_len and
            int(rec['full_len']) == o.full_len)
