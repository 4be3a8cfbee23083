"""
Synthetic workload of the shape BASELINE.json names: ONT-like reads (log-normal lengths, N50 about 10 kb) with
adapters planted at the ends (and now and then in the middle), and an adapter set with the SHAPE of Porechop's
built-in database (reference porechop_abi/adapters.py:77-464: 119 start + 117 end sequences, lengths 22-50,
5,747 bases in total).  The sequences themselves are seeded random ACGT -- there is no network for real data and
the reference's database is not copied; only its length histogram is reproduced here.
"""
import numpy as np

# length histograms of the reference's database (measured by importing porechop_abi/adapters.py)
START_LENGTHS = [22] * 2 + [23] + [24] * 108 + [25, 27, 28, 28, 33, 37, 38, 50]
END_LENGTHS = [22] * 4 + [23] + [24] * 109 + [27, 30, 38]
_LETTERS = np.frombuffer(b'ACGT', dtype=np.uint8)


def adapter_set(seed=7):
    """-> (list of adapter byte strings, start_ids, end_ids). 236 sequences, 5,747 bases."""
    rng = np.random.default_rng(seed)
    seqs = [_LETTERS[rng.integers(0, 4, n)].tobytes() for n in START_LENGTHS + END_LENGTHS]
    start_ids = list(range(len(START_LENGTHS)))
    end_ids = list(range(len(START_LENGTHS), len(seqs)))
    return seqs, start_ids, end_ids


def read_lengths(n_reads, rng, mean_log=8.75, sigma=0.75, lo=200, hi=200000):
    """log-normal lengths; with these defaults the length-weighted median (N50) is about 10 kb."""
    return np.clip(rng.lognormal(mean_log, sigma, n_reads), lo, hi).astype(np.int64)


def _mutate(rng, seq, rate):
    out = bytearray()
    for ch in seq:
        x = rng.random()
        if x < rate / 3:
            continue
        if x < 2 * rate / 3:
            out.append(int(_LETTERS[rng.integers(0, 4)])); out.append(ch); continue
        if x < rate:
            out.append(int(_LETTERS[rng.integers(0, 4)])); continue
        out.append(ch)
    return bytes(out)


def make_reads(n_reads, adapters, start_ids, end_ids, seed=1, lengths=None, end_rate=0.7, middle_rate=0.01,
               error=0.1):
    """
    -> (blob uint8 [total], offsets int64 [n_reads + 1]).  One contiguous ASCII blob; read r is
    blob[offsets[r]:offsets[r + 1]].  A fraction of the reads carries a (mutated, possibly truncated) start and/or
    end adapter of set 0; a few carry one in the middle.
    """
    rng = np.random.default_rng(seed)
    if lengths is None:
        lengths = read_lengths(n_reads, rng)
    lengths = np.asarray(lengths, dtype=np.int64)
    off = np.zeros(n_reads + 1, dtype=np.int64)
    np.cumsum(lengths, out=off[1:])
    blob = _LETTERS[rng.integers(0, 4, int(off[-1]), dtype=np.uint8)]
    a_start, a_end = adapters[start_ids[0]], adapters[end_ids[0]]
    plant = rng.random((n_reads, 3))
    for r in np.nonzero(plant[:, 0] < end_rate)[0]:
        s = _mutate(rng, a_start, error)[int(rng.integers(0, 6)):]
        k = min(len(s), int(lengths[r]))
        blob[off[r]:off[r] + k] = np.frombuffer(s[:k], dtype=np.uint8)
    for r in np.nonzero(plant[:, 1] < end_rate)[0]:
        s = _mutate(rng, a_end, error)
        s = s[:len(s) - int(rng.integers(0, 6))]
        k = min(len(s), int(lengths[r]))
        if k:
            blob[off[r + 1] - k:off[r + 1]] = np.frombuffer(s[len(s) - k:], dtype=np.uint8)
    for r in np.nonzero(plant[:, 2] < middle_rate)[0]:
        if lengths[r] > 1000:
            s = _mutate(rng, a_start, error)
            p = int(rng.integers(300, lengths[r] - 300 - len(s)))
            blob[off[r] + p:off[r] + p + len(s)] = np.frombuffer(s, dtype=np.uint8)
    return blob, off


def end_windows(blob, off, end_size=150, stride=160):
    """
    The two windows the reference aligns per read (nanopore_read.py:164,170,181,203: seq[:end_size] and
    seq[-end_size:]) as fixed-stride records.  -> (start_buf [n, stride], end_buf [n, stride], win_len int32 [n])
    """
    n = len(off) - 1
    lens = np.minimum(off[1:] - off[:-1], end_size).astype(np.int32)
    idx = np.arange(stride, dtype=np.int64)[None, :]
    valid = idx < lens[:, None]
    s_idx = np.minimum(off[:-1, None] + idx, len(blob) - 1)
    e_idx = np.clip(off[1:, None] - lens[:, None] + idx, 0, len(blob) - 1)
    start = np.where(valid, blob[s_idx], 0).astype(np.uint8)
    end = np.where(valid, blob[e_idx], 0).astype(np.uint8)
    return start, end, lens


class AdapterSet(object):
    """The attributes of the reference's Adapter (adapters.py:18-53) that the batched phases touch: a named set with an
    optional start and end sequence, and the best scores phase A leaves on it."""

    def __init__(self, name, start_sequence=None, end_sequence=None):
        self.name = name
        self.start_sequence = start_sequence        # (name, sequence) or None
        self.end_sequence = end_sequence
        self.best_start_score, self.best_end_score = 0.0, 0.0

    def best_start_or_end_score(self):
        return max(self.best_start_score, self.best_end_score)

    def is_barcode(self):
        return self.name.startswith('Barcode ')

    def barcode_direction(self):
        return 'reverse' if '_rev' in self.start_sequence[0] else 'forward'

    def get_barcode_name(self):
        names = [self.name] + ([self.start_sequence[0]] if self.start_sequence else []) + ([self.end_sequence[0]] if self.end_sequence else [])
        return sorted(names, key=lambda x: len(x))[0].replace(' ', '_')


def adapter_sets(adapters, start_ids, end_ids):
    """The synthetic sequences as adapter sets (set k = k-th start + k-th end sequence), for the phase functions."""
    sets = []
    for k in range(max(len(start_ids), len(end_ids))):
        s = ('set%d_start' % k, adapters[start_ids[k]].decode()) if k < len(start_ids) else None
        e = ('set%d_end' % k, adapters[end_ids[k]].decode()) if k < len(end_ids) else None
        sets.append(AdapterSet('set %d' % k, s, e))
    return sets
