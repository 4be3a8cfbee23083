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


# ---- check_compatibility (porechop_abi/ab_initio_src/compatibility.cpp), restated on top of the oracle's alignment path ----
_DNA = {ord(c): c.upper() for c in 'ACGTacgt'}
_DNA.update({ord('U'): 'T', ord('u'): 'T'})


def oracle_compat(seq1, seq2):
    """The flag get_flag computes (compatibility.cpp:121-156): String<Dna> conversion, longer sequence first
    (make_stringSet, :19-31), semi-global alignment with linear gaps 2 / -1 / -1 (compatibility.h:9-12), then the
    reference's own scans over the two gapped rows (:37-117), with its integer identity and its [start, end) error count."""
    a = ''.join(_DNA.get(ord(c), 'A') for c in seq1)
    b = ''.join(_DNA.get(ord(c), 'A') for c in seq2)
    s1, s2 = (b, a) if len(a) < len(b) else (a, b)
    ops, _, hh, hv, th, tv, _ = oracle_path(s1, s2, (2, -1, -1, -1))
    row1, row2 = [], []
    i = j = 0
    for _ in range(hh):
        row1.append(s1[i]); row2.append('-'); i += 1
    for _ in range(hv):
        row1.append('-'); row2.append(s2[j]); j += 1
    for op in ops.decode():
        if op == 'D':
            row1.append(s1[i]); row2.append(s2[j]); i += 1; j += 1
        elif op == 'V':
            row1.append('-'); row2.append(s2[j]); j += 1
        else:
            row1.append(s1[i]); row2.append('-'); i += 1
    for _ in range(th):
        row1.append(s1[i]); row2.append('-'); i += 1
    for _ in range(tv):
        row1.append('-'); row2.append(s2[j]); j += 1
    assert i == len(s1) and j == len(s2)
    n = len(row1)
    start = end = -1
    s1_started = s2_started = False
    for k in range(n):
        s1_started |= row1[k] != '-'; s2_started |= row2[k] != '-'
        if s1_started and s2_started:
            start = k; break
    e1 = e2 = False
    for k in range(n - 1, -1, -1):
        e1 |= row1[k] != '-'; e2 |= row2[k] != '-'
        if e1 and e2:
            end = k; break
    mapped = end - start + 1
    dist = sum(1 for k in range(start, end) if row1[k] != row2[k])
    identity = float(int((mapped - dist) * 100 / mapped)) if (mapped - dist) * 100 >= 0 else float(-(-(mapped - dist) * 100 // mapped))
    s1_start = sum(1 for k in range(start) if row1[k] != '-')          # toSourcePosition(row1, start)
    s1_end = sum(1 for k in range(end) if row1[k] != '-')
    included = s1_start > 3 or (len(s1) - s1_end) > 3
    flag = 1 if identity >= 87.5 else 0
    if included and flag == 1:
        flag = 2
    return flag


COMPAT_REF_SO = os.path.join(ORACLE_DIR, '_ref', 'compatibility_ref.so')
_compat_ref = None


def reference_compat(seq1, seq2):
    """The reference's own check_compatibility, compiled from its source into oracle/_ref (oracle/Makefile)."""
    global _compat_ref
    if _compat_ref is None:
        _compat_ref = ctypes.CDLL(COMPAT_REF_SO)
        _compat_ref.check_compatibility.argtypes = [ctypes.c_char_p, ctypes.c_char_p]
        _compat_ref.check_compatibility.restype = ctypes.c_int
    return int(_compat_ref.check_compatibility(_b(seq1), _b(seq2)))
