"""
Reads as one byte buffer plus per-read segments, for the batched phases (porechop_batch.py).

The reference loads every record into Python strings (misc.py:109-168: load_fasta_or_fastq -> load_fastq / load_fasta) and
NanoporeRead keeps seq.upper() (nanopore_read.py:32-44).  The alignment path only needs the bases: the C ABI takes a blob
and (offset, length) segments, and the device-side encoder maps a/c/g/t/u in either case like SeqAn's own table
(alphabet_residue_tabs.h:113-140), so the bytes of the file can go to the GPU as they are.  Joining and encoding a
million Python strings per phase costs more than the alignments (scripts/exp/batch_phases.py); a ReadBlob does that work
once, with numpy, at load time.

FASTQ: four-line records as the reference parses them (line.strip() of lines 4k .. 4k+3).  FASTA: sequences may span
lines, so they are compacted into a fresh buffer.  gzip is read through gzip.open, as in the reference.
"""
import gzip

import numpy as np

_WS = (9, 10, 11, 12, 13, 32)                     # what str.strip() removes from ASCII text


def _open_bytes(filename):
    with open(filename, 'rb') as f:
        magic = f.read(3)
    opener = gzip.open if magic == b'\x1f\x8b\x08' else open          # misc.py:60-81
    with opener(filename, 'rb') as f:
        return f.read()


def _strip(buf, start, end):
    """Vectorised line.strip(): move start / end inwards over ASCII whitespace."""
    start, end = start.copy(), end.copy()
    ws = np.zeros(256, dtype=bool)
    ws[list(_WS)] = True
    while True:
        m = (end > start) & ws[buf[np.maximum(end - 1, 0)]]
        if not m.any():
            break
        end[m] -= 1
    while True:
        m = (end > start) & ws[buf[np.minimum(start, len(buf) - 1)]]
        if not m.any():
            break
        start[m] += 1
    return start, end


class ReadBlob(object):
    """buf: uint8 buffer; read i is buf[seq_off[i] : seq_off[i] + seq_len[i]]; names / qualities likewise (FASTQ)."""

    def __init__(self, buf, seq_off, seq_len, name_off, name_len, qual_off=None, qual_len=None, file_type='FASTQ'):
        self.buf = buf
        self.seq_off = np.ascontiguousarray(seq_off, dtype=np.int64)
        self.seq_len = np.ascontiguousarray(seq_len, dtype=np.int64)
        self.name_off, self.name_len = name_off, name_len
        self.qual_off, self.qual_len = qual_off, qual_len
        self.file_type = file_type
        # per-read state of the phases, as arrays: the BlobRead objects of make_reads() read and write these, and the
        # batched phases update them with numpy instead of touching a Python object per hit
        self.start_trim = np.zeros(len(self.seq_off), dtype=np.int64)
        self.end_trim = np.zeros(len(self.seq_off), dtype=np.int64)
        self.end_hits = {False: None, True: None}      # at_end -> (first hit of every read [n + 1], sets, k, full, partial, start, end)

    def __len__(self):
        return len(self.seq_off)

    @classmethod
    def from_buffer(cls, buf, offsets):
        """Reads that already sit back to back in one buffer: read i = buf[offsets[i] : offsets[i + 1]]; names r0, r1, ..."""
        offsets = np.asarray(offsets, dtype=np.int64)
        n = len(offsets) - 1
        z = np.zeros(n, dtype=np.int64)
        return cls(np.asarray(buf, dtype=np.uint8), offsets[:-1], offsets[1:] - offsets[:-1], z, z, None, None, 'FASTA')

    def head(self, k):
        """The first k reads (views, nothing is copied)."""
        q = (None, None) if self.qual_off is None else (self.qual_off[:k], self.qual_len[:k])
        return ReadBlob(self.buf, self.seq_off[:k], self.seq_len[:k], self.name_off[:k], self.name_len[:k], q[0], q[1], self.file_type)

    def take(self, idx):
        """The reads idx (ascending), as views of the same buffer: e.g. the check reads of a directory of files."""
        idx = np.asarray(idx, dtype=np.int64)
        q = (None, None) if self.qual_off is None else (self.qual_off[idx], self.qual_len[idx])
        return ReadBlob(self.buf, self.seq_off[idx], self.seq_len[idx], np.asarray(self.name_off)[idx], np.asarray(self.name_len)[idx],
                        q[0], q[1], self.file_type)

    def compact(self):
        """The same reads in a buffer of their own that holds nothing else (a rank's shard of a file: what it sends to its GPU)."""
        def gather(off, ln):
            off, ln = np.asarray(off, dtype=np.int64), np.asarray(ln, dtype=np.int64)
            new_off = np.zeros(len(ln) + 1, dtype=np.int64)
            np.cumsum(ln, out=new_off[1:])
            pos = np.repeat(off - new_off[:-1], ln) + np.arange(int(new_off[-1]), dtype=np.int64)
            return self.buf[pos], new_off[:-1]
        seq, seq_off = gather(self.seq_off, self.seq_len)
        name, name_off = gather(self.name_off, self.name_len)
        parts, q_off = [seq, name], None
        if self.qual_off is not None:
            qual, q_off = gather(self.qual_off, self.qual_len)
            parts.append(qual)
            q_off = q_off + len(seq) + len(name)
        return ReadBlob(np.concatenate(parts), seq_off, self.seq_len, name_off + len(seq), np.asarray(self.name_len, dtype=np.int64),
                        q_off, None if q_off is None else np.asarray(self.qual_len, dtype=np.int64), self.file_type)

    @classmethod
    def concat(cls, blobs):
        """The reads of several blobs (the FASTQ files of a directory) behind one another, as ONE buffer."""
        blobs = list(blobs)
        if len(blobs) == 1:
            return blobs[0]
        shift = np.cumsum([0] + [len(b.buf) for b in blobs[:-1]])
        cat = lambda get: np.concatenate([np.asarray(get(b), dtype=np.int64) for b in blobs])
        cat_off = lambda get: np.concatenate([np.asarray(get(b), dtype=np.int64) + s for b, s in zip(blobs, shift)])
        has_q = all(b.qual_off is not None for b in blobs)
        return cls(np.concatenate([b.buf for b in blobs]), cat_off(lambda b: b.seq_off), cat(lambda b: b.seq_len),
                   cat_off(lambda b: b.name_off), cat(lambda b: b.name_len),
                   cat_off(lambda b: b.qual_off) if has_q else None, cat(lambda b: b.qual_len) if has_q else None, blobs[0].file_type)

    def owns(self, reads):
        """Are these the BlobRead objects of this blob, in order (then their state is this blob's arrays)?"""
        return len(reads) == len(self) and len(reads) > 0 and getattr(reads[0], '_blob', None) is self and getattr(reads[-1], '_blob', None) is self

    def apply_end_hits(self, at_end, sets, read_idx, set_idx, full, partial, r_start, r_end, end_size, extra_trim_size):
        """find_start_trim / find_end_trim for all reads at once (nanopore_read.py:187-195, :209-217): the kept alignments
        (ordered by read, then by set) become the reads' start / end alignment lists, the trim amounts their maxima."""
        n = len(self)
        amount = (end_size - r_start) + extra_trim_size if at_end else r_end + extra_trim_size
        trim = self.end_trim if at_end else self.start_trim
        if len(read_idx):
            first = np.flatnonzero(np.r_[True, read_idx[1:] != read_idx[:-1]])
            reads_hit = read_idx[first]
            trim[reads_hit] = np.maximum(trim[reads_hit], np.maximum.reduceat(amount.astype(np.int64), first))
        bounds = np.searchsorted(read_idx, np.arange(n + 1))
        self.end_hits[at_end] = (bounds, list(sets), set_idx, full, partial, r_start, r_end)

    def alignments(self, i, at_end):
        """The (adapter set, full score, partial score, read start, read end) tuples of read i, as the reference appends them."""
        h = self.end_hits[at_end]
        if h is None:
            return []
        bounds, sets, k, full, partial, rs, re = h
        a, b = int(bounds[i]), int(bounds[i + 1])
        return [(sets[int(k[t])], float(full[t]), float(partial[t]), int(rs[t]), int(re[t])) for t in range(a, b)]

    def make_reads(self):
        """One BlobRead per record: what the phase functions need of the reference's NanoporeRead, the sequence on demand."""
        return [BlobRead(self, i) for i in range(len(self))]

    def seq(self, i):
        """The sequence as NanoporeRead stores it (nanopore_read.py:36-39): upper case, U -> T when U outnumbers T."""
        seg = self.buf[self.seq_off[i]:self.seq_off[i] + self.seq_len[i]]
        raw = seg.tobytes()
        s = raw.decode()
        if len(seg) and seg.max() > 96:                     # a lower-case letter somewhere (str.isupper costs 3 x str.upper)
            s = s.upper()
        if b'U' in raw or b'u' in raw:                       # (a memchr: counting is only needed when there is a U at all)
            if s.count('U') > s.count('T'):
                s = s.replace('U', 'T')
        return s

    def is_rna(self, i):
        raw = self.buf[self.seq_off[i]:self.seq_off[i] + self.seq_len[i]].tobytes()
        if b'U' not in raw and b'u' not in raw:
            return False
        s = raw.decode().upper()
        return s.count('U') > s.count('T')

    def full_name(self, i):
        return self.buf[self.name_off[i]:self.name_off[i] + self.name_len[i]].tobytes().decode()

    def name(self, i):
        return self.full_name(i).split()[0]

    def quals(self, i):
        if self.qual_off is None:
            return ''
        return self.buf[self.qual_off[i]:self.qual_off[i] + self.qual_len[i]].tobytes().decode()

    def windows(self, end_size, at_end):
        """seq[:end_size] / seq[-end_size:] of every read (nanopore_read.py:164,170,181,203) as a fixed-stride buffer."""
        stride = max(16, -(-end_size // 16) * 16)
        lens = np.minimum(self.seq_len, end_size)
        first = self.seq_off + (self.seq_len - lens if at_end else 0)
        out = np.zeros((len(self), stride), dtype=np.uint8)
        # rows of `stride` contiguous bytes out of a sliding-window view (a row copy each, no index matrix); the few windows
        # that start within `stride` bytes of the buffer's end are copied one by one
        fits = first <= len(self.buf) - stride
        if len(self.buf) >= stride and fits.any():
            out[fits] = np.lib.stride_tricks.sliding_window_view(self.buf, stride)[first[fits]]
        for i in np.flatnonzero(~fits):
            out[i, :lens[i]] = self.buf[first[i]:first[i] + lens[i]]
        out[np.arange(stride, dtype=np.int32)[None, :] >= lens[:, None]] = 0
        return np.ascontiguousarray(out, dtype=np.uint8), lens.astype(np.int32)

    def trimmed_segments(self, start_trim, end_trim):
        """(offset, length) in buf of seq[start_trim : len - end_trim] for every read, with Python's slice rules (a stop
        index below zero counts from the end), and the whole read where both trims are 0 (nanopore_read.py:77-81)."""
        n = self.seq_len
        start_trim = np.asarray(start_trim, dtype=np.int64)
        end_trim = np.asarray(end_trim, dtype=np.int64)
        lo = np.minimum(start_trim, n)
        stop = n - end_trim
        stop = np.where(stop < 0, np.maximum(stop + n, 0), np.minimum(stop, n))
        hi = np.maximum(lo, stop)
        return self.seq_off + lo, hi - lo


class BlobRead(object):
    """A read of a ReadBlob with the attributes of the reference's NanoporeRead that the three phases touch
    (nanopore_read.py:32-75).  Building one costs two assignments: the per-read sets, dicts and strings NanoporeRead.__init__
    creates come into being when someone first looks at them (most reads never have a middle hit or a barcode score), and
    the sequence string is only built when someone asks for it (a read with a middle hit, the writer)."""
    __slots__ = ('_blob', '_i', '_seq', '_start_al', '_end_al',
                 'middle_adapter_positions', 'middle_trim_positions', 'middle_hit_str', 'start_barcode_scores', 'end_barcode_scores',
                 'barcode_call', 'albacore_barcode_call', 'best_start_barcode', 'second_best_start_barcode', 'best_end_barcode',
                 'second_best_end_barcode')
    _FRESH = {'_seq': lambda: None, '_start_al': lambda: None, '_end_al': lambda: None,
              'middle_adapter_positions': set, 'middle_trim_positions': set, 'middle_hit_str': str,
              'start_barcode_scores': dict, 'end_barcode_scores': dict,
              'barcode_call': lambda: 'none', 'albacore_barcode_call': lambda: None,
              'best_start_barcode': lambda: ('none', 0.0), 'second_best_start_barcode': lambda: ('none', 0.0),
              'best_end_barcode': lambda: ('none', 0.0), 'second_best_end_barcode': lambda: ('none', 0.0)}

    def __init__(self, blob, i):
        self._blob, self._i = blob, i

    def __getattr__(self, name):
        # only reached for a slot nobody has written yet: give it the value NanoporeRead.__init__ would have given it
        fresh = BlobRead._FRESH.get(name)
        if fresh is None:
            raise AttributeError(name)
        value = fresh()
        object.__setattr__(self, name, value)
        return value

    @property
    def name(self):
        """The whole header line, as NanoporeRead.name holds it (porechop_abi.py:143-146: x[4] / x[2] of the loader's tuples)."""
        return self._blob.full_name(self._i) if self._blob.name_len[self._i] else 'r%d' % self._i

    # trims live in the blob's arrays; the alignment lists are built from the blob's hit arrays when someone looks at them
    @property
    def start_trim_amount(self):
        return int(self._blob.start_trim[self._i])

    @start_trim_amount.setter
    def start_trim_amount(self, v):
        self._blob.start_trim[self._i] = v

    @property
    def end_trim_amount(self):
        return int(self._blob.end_trim[self._i])

    @end_trim_amount.setter
    def end_trim_amount(self, v):
        self._blob.end_trim[self._i] = v

    @property
    def start_adapter_alignments(self):
        if self._start_al is None:
            self._start_al = self._blob.alignments(self._i, False)
        return self._start_al

    @property
    def end_adapter_alignments(self):
        if self._end_al is None:
            self._end_al = self._blob.alignments(self._i, True)
        return self._end_al

    @property
    def seq(self):
        if self._seq is None:
            self._seq = self._blob.seq(self._i)
        return self._seq

    @property
    def quals(self):
        """Quality string, padded with '+' to the sequence's length as NanoporeRead pads it (nanopore_read.py:44-46)."""
        q = self._blob.quals(self._i)
        n = int(self._blob.seq_len[self._i])
        return q + '+' * (n - len(q)) if len(q) < n else q

    @property
    def rna(self):
        """More U than T in the sequence as loaded (nanopore_read.py:36-41): output turns T back into U."""
        return self._blob.is_rna(self._i)

    def get_seq_with_start_end_adapters_trimmed(self):
        if not self.start_trim_amount and not self.end_trim_amount:
            return self.seq
        return self.seq[self.start_trim_amount:len(self.seq) - self.end_trim_amount]

    def adapters_found(self):
        """The fork's filter (nanopore_read.py:23-30, porechop_abi.py:36-39): an adapter at BOTH ends."""
        return bool(self.start_adapter_alignments) and bool(self.end_adapter_alignments)

    # what the reference's output_reads (porechop_abi.py:535-668) calls on a read: with these a BlobRead list goes through the
    # reference's own writer unchanged (read_output.output_reads is the same writer for this repo's callers)
    def seq_length_with_start_end_adapters_trimmed(self):
        return len(self.get_seq_with_start_end_adapters_trimmed())

    def get_split_read_parts(self, min_split_read_size):
        from .read_output import get_split_read_parts
        return get_split_read_parts(self, min_split_read_size)

    def get_fasta(self, min_split_read_size, discard_middle, untrimmed=False):
        from .read_output import get_fasta
        return get_fasta(self, min_split_read_size, discard_middle, untrimmed)

    def get_fastq(self, min_split_read_size, discard_middle, untrimmed=False):
        from .read_output import get_fastq
        return get_fastq(self, min_split_read_size, discard_middle, untrimmed)


def load_read_blob(filename):
    """The records of a FASTQ or FASTA file (plain or gzip) as a ReadBlob; record order = file order."""
    raw = _open_bytes(filename)
    buf = np.frombuffer(raw, dtype=np.uint8)
    if len(buf) == 0:
        raise ValueError('File is neither FASTA or FASTQ')
    nl = np.flatnonzero(buf == 10)
    starts = np.concatenate(([0], nl + 1))
    ends = np.concatenate((nl, [len(buf)]))
    if starts[-1] >= len(buf):                               # the file ends with a newline: no line after it
        starts, ends = starts[:-1], ends[:-1]
    first = chr(buf[0])
    if first == '@':
        n = len(starts) // 4                                 # a truncated last record: the reference stops with an error there
        if len(starts) % 4:
            tail_s, tail_e = _strip(buf, starts[4 * n:], ends[4 * n:])
            if (tail_e > tail_s).any():
                raise IndexError('FASTQ record count is not a multiple of four lines')
        s, e = _strip(buf, starts[:4 * n], ends[:4 * n])
        s, e = s.reshape(n, 4), e.reshape(n, 4)
        name_off = np.minimum(s[:, 0] + 1, e[:, 0])          # line.strip()[1:]
        return ReadBlob(buf, s[:, 1], e[:, 1] - s[:, 1], name_off, e[:, 0] - name_off, s[:, 3], e[:, 3] - s[:, 3], 'FASTQ')
    if first == '>':
        s, e = _strip(buf, starts, ends)
        keep = e > s                                          # blank lines are skipped (misc.py:136-138)
        s, e = s[keep], e[keep]
        header = buf[s] == ord('>')
        rec = np.cumsum(header) - 1                           # record of every line
        n = int(header.sum())
        body = ~header & (rec >= 0)
        line_len = (e - s)[body]
        seq_len = np.bincount(rec[body], weights=line_len, minlength=n).astype(np.int64)
        seq_off = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(seq_len, out=seq_off[1:])
        # compact the sequence lines: one gather over the bytes that belong to them
        pos = np.repeat(s[body] - np.concatenate(([0], np.cumsum(line_len)[:-1])), line_len) + np.arange(int(line_len.sum()), dtype=np.int64)
        names = buf[np.repeat(s[header] + 1, (e - s)[header] - 1) + (np.arange(int(((e - s)[header] - 1).sum()), dtype=np.int64)
                                                                    - np.repeat(np.concatenate(([0], np.cumsum((e - s)[header] - 1)[:-1])), (e - s)[header] - 1))]
        name_len = (e - s)[header] - 1
        name_off = np.concatenate(([0], np.cumsum(name_len)[:-1])) + int(seq_off[-1])
        out = np.concatenate((buf[pos], names)) if n else np.zeros(0, dtype=np.uint8)
        return ReadBlob(out, seq_off[:-1], seq_len, name_off, name_len, None, None, 'FASTA')
    raise ValueError('File is neither FASTA or FASTQ')


def get_albacore_barcode_from_path(albacore_path):
    """Albacore's bin of a file, from the directory it lies in (porechop_abi.py:190-197)."""
    import re
    if '/unclassified/' in albacore_path:
        return 'none'
    matches = re.findall('/barcode(\\d\\d)/', albacore_path)
    return 'BC' + matches[-1] if matches else None


def load_reads(input_file_or_directory, check_read_count):
    """
    porechop_abi.py:133-187 on a ReadBlob: a file, or a directory searched recursively for .fastq / .fastq.gz files (an
    Albacore output directory: every read gets its file's bin as albacore_barcode_call, the check reads are the first
    round(check_read_count / files) reads of every file).
    -> (blob, reads = its BlobRead objects, check_reads, check_blob, read_type)
    """
    import os
    import sys
    if os.path.isfile(input_file_or_directory):
        blob = load_read_blob(input_file_or_directory)
        reads = blob.make_reads()
        k = min(check_read_count, len(reads))
        return blob, reads, reads[:k], blob.head(k), blob.file_type
    if os.path.isdir(input_file_or_directory):
        fastqs = sorted(os.path.join(dir_path, f) for dir_path, _, filenames in os.walk(input_file_or_directory)
                        for f in filenames if f.lower().endswith('.fastq') or f.lower().endswith('.fastq.gz'))
        if not fastqs:
            sys.exit('Error: could not find fastq files in ' + input_file_or_directory)
        per_file = int(round(check_read_count / len(fastqs)))
        blobs = [load_read_blob(f) for f in fastqs]
        blob = ReadBlob.concat(blobs)
        reads = blob.make_reads()
        check_idx, first = [], 0
        for f, b in zip(fastqs, blobs):
            call = get_albacore_barcode_from_path(f)
            for r in reads[first:first + len(b)]:
                r.albacore_barcode_call = call
            check_idx.extend(range(first, first + min(per_file, len(b))))
            first += len(b)
        return blob, reads, [reads[i] for i in check_idx], blob.take(check_idx), 'FASTQ'
    sys.exit('Error: could not find ' + input_file_or_directory)
