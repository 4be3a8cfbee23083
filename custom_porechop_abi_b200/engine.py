"""
AlignmentEngine -- batched submit/collect over the C ABI (include/porechop_b200.h).

This is the host-side replacement for the reference's per-pair call inside a thread pool
(porechop_abi.py:228-239, :406-432, :500-517 -> nanopore_read.align_adapter, nanopore_read.py:485-500):
one engine per GPU, adapters and scoring scheme set once, then whole phases submitted at a time.
Results come back as a numpy structured array with the integer fields the reference's result string is
printed from; ``identities()`` evaluates the two percent identities in double exactly as
src/alignment.cpp:81,89 does, and ``to_align_adapter_tuple`` reproduces what nanopore_read.align_adapter
returns from the string.
"""
import contextlib
import ctypes
from ctypes import byref, c_float, c_void_p

import numpy as np

from . import cpp_function_wrappers as w

RESULT_DTYPE = np.dtype([('read_start', '<i4'), ('read_end', '<i4'), ('adapter_start', '<i4'),
                         ('adapter_end', '<i4'), ('raw_score', '<i4'), ('matches', '<i4'),
                         ('aligned_len', '<i4'), ('full_len', '<i4')])
assert RESULT_DTYPE.itemsize == ctypes.sizeof(w.PcbResult)
BEST_DTYPE = np.dtype([('matches', '<i4'), ('full_len', '<i4')])
HIT_DTYPE = np.dtype([('window', '<i4'), ('id_pos', '<i4'), ('read_start', '<i4'), ('read_end', '<i4'),
                      ('raw_score', '<i4'), ('matches', '<i4'), ('aligned_len', '<i4'), ('full_len', '<i4')])
TOP2_DTYPE = np.dtype([('best', '<i2'), ('second', '<i2'), ('best_ident', '<u2'), ('second_ident', '<u2')])
assert BEST_DTYPE.itemsize == ctypes.sizeof(w.PcbBest) and HIT_DTYPE.itemsize == ctypes.sizeof(w.PcbHit)


def threshold_e6(threshold):
    """floor(threshold * 10^6), computed in decimal (75.3 * 1e6 is not an integer in binary floating point)."""
    import decimal
    return int((decimal.Decimal(repr(float(threshold))) * 1000000).to_integral_value(rounding=decimal.ROUND_FLOOR))


class WindowHits(object):
    """What pcb_align_windows_hits returns: `hits` (HIT_DTYPE, ordered by window, then by adapter position) and `ident`
    ([n_windows, n_wanted] uint16 = matches << 9 | full_len, 0 = failed alignment).  Valid once the call is collected
    (at once, or after AlignmentEngine.wait() in deferred mode)."""

    def __init__(self, hits_buf, n_hits, ident):
        self._buf, self._n, self.ident = hits_buf, n_hits, ident

    @property
    def n_hits(self):
        return int(self._n.value)

    @property
    def hits(self):
        return self._buf[:min(self.n_hits, len(self._buf))]

    def identities(self):
        """full-adapter percent identity of every (window, wanted adapter) pair as the reference parses it from "%f"."""
        m = (self.ident >> 9).astype(np.int64)
        length = (self.ident & 0x1FF).astype(np.int64)
        return m, length



def _ptr(a):
    return a.ctypes.data_as(c_void_p)


class AlignmentEngine(object):
    def __init__(self, device=0, scoring_scheme_vals=(3, -6, -5, -2)):
        self._h = c_void_p()
        rc = w.C_LIB.pcb_create(int(device), byref(self._h))
        if rc != 0:
            raise w.AlignmentError('pcb_create failed (%d): %s' % (rc, w.C_LIB.pcb_last_error(None).decode()))
        self.device = device
        self.adapters = []
        self._slot = 0
        self._shapes = {}
        self.set_scoring(scoring_scheme_vals)

    # -- plumbing -----------------------------------------------------------------------------
    def _check(self, rc, what):
        if rc != 0:
            raise w.AlignmentError('%s failed (%d): %s' % (what, rc, w.C_LIB.pcb_last_error(self._h).decode()))

    def close(self):
        if self._h:
            w.C_LIB.pcb_destroy(self._h)
            self._h = c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- configuration ------------------------------------------------------------------------
    def set_scoring(self, vals):
        """[match, mismatch, gap_open, gap_extend], the order of --scoring_scheme (arg_parser.py:178,230-236)."""
        self.scoring = tuple(int(v) for v in vals)
        self._check(w.C_LIB.pcb_set_scoring(self._h, *self.scoring), 'pcb_set_scoring')

    def select_slot(self, slot):
        """Pick one of the 8 staging slots (inputs + results) the following stage/run/fetch calls act on."""
        self._check(w.C_LIB.pcb_select_slot(self._h, int(slot)), 'pcb_select_slot')
        self._slot = int(slot)

    def set_adapters(self, sequences):
        """sequences: list of str / bytes. Adapter ids used everywhere else are indices into this list."""
        seqs = [s.encode() if isinstance(s, str) else bytes(s) for s in sequences]
        blob = b''.join(seqs)
        off = np.zeros(len(seqs) + 1, dtype=np.int64)
        if seqs:
            off[1:] = np.cumsum([len(s) for s in seqs])
        self._check(w.C_LIB.pcb_set_adapters(self._h, blob, off.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                             len(seqs)), 'pcb_set_adapters')
        self.adapters = seqs

    # -- windows x adapters (phases A and B) --------------------------------------------------------
    @staticmethod
    def pack_windows(windows, stride=None):
        """list of str/bytes windows -> (uint8 array [n, stride], int32 lengths)."""
        ws = [x.encode() if isinstance(x, str) else bytes(x) for x in windows]
        lens = np.array([len(x) for x in ws], dtype=np.int32)
        if stride is None:
            stride = max(16, int(-(-max([1] + [len(x) for x in ws]) // 16) * 16))
        buf = np.zeros((len(ws), stride), dtype=np.uint8)
        for i, x in enumerate(ws):
            buf[i, :len(x)] = np.frombuffer(x, dtype=np.uint8)
        return buf, lens

    def align_windows(self, win_buf, win_len, adapter_ids):
        """Every window against every listed adapter. Returns records of shape [n_windows, len(adapter_ids)]."""
        win_buf = np.ascontiguousarray(win_buf, dtype=np.uint8)
        win_len = np.ascontiguousarray(win_len, dtype=np.int32)
        ids = np.ascontiguousarray(adapter_ids, dtype=np.int32)
        n, stride = win_buf.shape
        out = np.empty((n, len(ids)), dtype=RESULT_DTYPE)
        self._check(w.C_LIB.pcb_align_windows(self._h, _ptr(win_buf), n, stride, _ptr(win_len), _ptr(ids), len(ids),
                                              _ptr(out)), 'pcb_align_windows')
        if getattr(self, '_deferred', False):
            self._held.append((win_buf, win_len, out))
        return out

    def _window_args(self, win_buf, win_len):
        """(buffer, lengths, n, stride) of a window argument; win_buf None = the slot's staged windows (stage_windows_from_reads),
        win_len then is their number."""
        if win_buf is None:
            return None, None, int(win_len), 0
        win_buf = np.ascontiguousarray(win_buf, dtype=np.uint8)
        return win_buf, np.ascontiguousarray(win_len, dtype=np.int32), win_buf.shape[0], win_buf.shape[1]

    def stage_windows_from_reads(self, read_off, read_len, end_size, at_end):
        """The end windows of the reads, cut out of the announced blob (upload_reads) on the device into the selected slot
        (pcb_stage_windows_from_reads).  Then align_windows_best / align_windows_hits(None, n_reads, ...)."""
        read_off = np.ascontiguousarray(read_off, dtype=np.int64)
        read_len = np.ascontiguousarray(read_len, dtype=np.int32)
        self._check(w.C_LIB.pcb_stage_windows_from_reads(self._h, _ptr(read_off), _ptr(read_len), len(read_off), int(end_size),
                                                         1 if at_end else 0), 'pcb_stage_windows_from_reads')
        if getattr(self, '_deferred', False):
            self._held.append((read_off, read_len))
        self._shapes[self._slot] = (len(read_off), None)
        return len(read_off)

    def align_windows_best(self, win_buf, win_len, adapter_ids):
        """Phase A in one call: per listed adapter the (matches, full_len) of its best full-adapter identity over all windows
        (pcb_align_windows_best).  Returns a BEST_DTYPE array of len(adapter_ids)."""
        win_buf, win_len, n, stride = self._window_args(win_buf, win_len)
        ids = np.ascontiguousarray(adapter_ids, dtype=np.int32)
        out = np.zeros(len(ids), dtype=BEST_DTYPE)
        self._check(w.C_LIB.pcb_align_windows_best(self._h, None if win_buf is None else _ptr(win_buf), n, stride,
                                                   None if win_len is None else _ptr(win_len), _ptr(ids), len(ids), _ptr(out)),
                    'pcb_align_windows_best')
        if getattr(self, '_deferred', False):
            self._held.append((win_buf, win_len, ids, out))
        return out

    def align_windows_hits(self, win_buf, win_len, adapter_ids, end_threshold, at_end, end_size, min_trim_size,
                           want_ident=None, hits_cap=None, hits_buf=None, ident_buf=None):
        """Phase B in one call (pcb_align_windows_hits): only the pairs find_start_trim / find_end_trim would keep, plus the
        full-adapter identity of every window for the adapters flagged in want_ident.  Returns a WindowHits."""
        win_buf, win_len, n, stride = self._window_args(win_buf, win_len)
        ids = np.ascontiguousarray(adapter_ids, dtype=np.int32)
        flt = w.PcbHitFilter(threshold_e6(end_threshold), 1 if at_end else 0, int(end_size), int(min_trim_size), 0)
        want = None if want_ident is None else np.ascontiguousarray(want_ident, dtype=np.uint8)
        n_want = 0 if want is None else int(np.count_nonzero(want))
        if hits_buf is None:
            cap = int(hits_cap) if hits_cap is not None else int(min(n * len(ids), max(16 * n, 4096)))
            hits_buf = np.empty(max(cap, 1), dtype=HIT_DTYPE)
        keep_ident = ident_buf is not False                      # ident_buf=False: the identities stay on the device (fetch_top2)
        ident = ident_buf if (ident_buf is not None and keep_ident) else np.zeros((n, n_want if keep_ident else 0), dtype=np.uint16)
        n_hits = ctypes.c_int64(0)
        deferred = getattr(self, '_deferred', False)
        wp, lp = (None if win_buf is None else _ptr(win_buf)), (None if win_len is None else _ptr(win_len))
        rc = w.C_LIB.pcb_align_windows_hits(self._h, wp, n, stride, lp, _ptr(ids), len(ids), byref(flt),
                                            None if want is None else _ptr(want), _ptr(hits_buf), len(hits_buf), byref(n_hits),
                                            _ptr(ident) if (n_want and keep_ident) else None)
        if rc == 3 and not deferred and n_hits.value > len(hits_buf):      # more hits than room: once more with room for all
            hits_buf = np.empty(int(n_hits.value), dtype=HIT_DTYPE)
            rc = w.C_LIB.pcb_align_windows_hits(self._h, wp, n, stride, lp, _ptr(ids), len(ids), byref(flt),
                                                None if want is None else _ptr(want), _ptr(hits_buf), len(hits_buf), byref(n_hits),
                                                _ptr(ident) if (n_want and keep_ident) else None)
        self._check(rc, 'pcb_align_windows_hits')
        res = WindowHits(hits_buf, n_hits, ident)
        if deferred:
            self._held.append((win_buf, win_len, ids, want, flt, res))
        return res

    def fetch_top2(self, n_windows, out=None):
        """The best and second-best barcode identity of every window of this slot's last align_windows_hits call
        (pcb_fetch_top2): TOP2_DTYPE records, positions counted among the wanted adapters."""
        out = np.empty(n_windows, dtype=TOP2_DTYPE) if out is None else out
        self._check(w.C_LIB.pcb_fetch_top2(self._h, _ptr(out), n_windows), 'pcb_fetch_top2')
        if getattr(self, '_deferred', False):
            self._held.append((out,))
        return out

    def fetch_trims(self, n_windows, out=None):
        """Per window of this slot's last align_windows_hits call the trim amount its kept alignments ask for, before
        extra_trim_size (pcb_fetch_trims): int32, -1 = none kept."""
        out = np.empty(n_windows, dtype=np.int32) if out is None else out
        self._check(w.C_LIB.pcb_fetch_trims(self._h, _ptr(out), n_windows), 'pcb_fetch_trims')
        if getattr(self, '_deferred', False):
            self._held.append((out,))
        return out

    def stage_windows(self, win_buf, win_len, adapter_ids):
        win_buf = np.ascontiguousarray(win_buf, dtype=np.uint8)
        win_len = np.ascontiguousarray(win_len, dtype=np.int32)
        ids = np.ascontiguousarray(adapter_ids, dtype=np.int32)
        n, stride = win_buf.shape
        self._check(w.C_LIB.pcb_stage_windows(self._h, _ptr(win_buf), n, stride, _ptr(win_len)), 'pcb_stage_windows')
        self._check(w.C_LIB.pcb_stage_ids(self._h, _ptr(ids), len(ids)), 'pcb_stage_ids')
        self.synchronize()
        self._shapes[self._slot] = (n, len(ids))

    def run_windows_resident(self):
        self._check(w.C_LIB.pcb_run_windows_resident(self._h), 'pcb_run_windows_resident')

    def fetch(self, shape=None):
        shape = shape or self._shapes[self._slot]
        out = np.empty(shape, dtype=RESULT_DTYPE)
        self._check(w.C_LIB.pcb_fetch_results(self._h, _ptr(out), int(np.prod(shape))), 'pcb_fetch_results')
        return out

    # -- explicit tasks (phase C and anything irregular) ------------------------------------------------
    def upload_reads(self, blob):
        """Announces the blob of a later align_tasks(None, ...): it travels to the GPU under the window calls made in
        between (pcb_upload_reads).  The array is kept alive here until that call."""
        blob = np.frombuffer(blob, dtype=np.uint8) if isinstance(blob, (bytes, bytearray)) else np.ascontiguousarray(blob, dtype=np.uint8)
        self._check(w.C_LIB.pcb_upload_reads(self._h, _ptr(blob), blob.size), 'pcb_upload_reads')
        self._uploaded = blob

    def align_tasks(self, blob, seg_off, seg_len, task_adapter, out=None):
        """task t = blob[seg_off[t] : seg_off[t] + seg_len[t]] vs adapter task_adapter[t]; any length.
        blob None: the blob announced with upload_reads.  out: result buffer to fill (e.g. pinned), else a new array."""
        seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        seg_len = np.ascontiguousarray(seg_len, dtype=np.int32)
        ad = np.ascontiguousarray(task_adapter, dtype=np.int32)
        if out is None:
            out = np.empty(len(ad), dtype=RESULT_DTYPE)
        if blob is None:
            if getattr(self, '_uploaded', None) is None:
                raise w.AlignmentError('align_tasks(None, ...) without upload_reads')
            held = self._uploaded                     # stays announced until the batch is collected: several task calls may use it
            ptr, size = None, held.size
        else:
            self._uploaded = None
            held = np.frombuffer(blob, dtype=np.uint8) if isinstance(blob, (bytes, bytearray)) else np.ascontiguousarray(blob, dtype=np.uint8)
            ptr, size = _ptr(held), held.size
        self._check(w.C_LIB.pcb_align_tasks(self._h, ptr, size, _ptr(seg_off), _ptr(seg_len), _ptr(ad), len(ad), _ptr(out)),
                    'pcb_align_tasks')
        if getattr(self, '_deferred', False):
            self._held.append((held, seg_off, seg_len, ad, out))
        return out

    def stage_tasks(self, blob, seg_off, seg_len, task_adapter):
        blob = np.frombuffer(blob, dtype=np.uint8) if isinstance(blob, (bytes, bytearray)) else np.ascontiguousarray(blob, dtype=np.uint8)
        seg_off = np.ascontiguousarray(seg_off, dtype=np.int64)
        seg_len = np.ascontiguousarray(seg_len, dtype=np.int32)
        ad = np.ascontiguousarray(task_adapter, dtype=np.int32)
        self._check(w.C_LIB.pcb_stage_tasks(self._h, _ptr(blob), blob.size, _ptr(seg_off), _ptr(seg_len), _ptr(ad),
                                            len(ad)), 'pcb_stage_tasks')
        self.synchronize()
        self._shapes[self._slot] = (len(ad),)

    def run_tasks_resident(self):
        self._check(w.C_LIB.pcb_run_tasks_resident(self._h), 'pcb_run_tasks_resident')

    def align_pairs(self, pairs):
        """Convenience: list of (read_str, adapter_id) -> records."""
        seqs = [p[0].encode() if isinstance(p[0], str) else bytes(p[0]) for p in pairs]
        lens = np.array([len(s) for s in seqs], dtype=np.int32)
        off = np.zeros(len(seqs), dtype=np.int64)
        if len(seqs) > 1:
            off[1:] = np.cumsum(lens[:-1], dtype=np.int64)
        return self.align_tasks(b''.join(seqs), off, lens, [p[1] for p in pairs])

    def check_compatibility(self, sequences, pairs):
        """flags (0 / 1 / 2, compatibility.cpp:121-156) for the index pairs of `sequences`, in one submit.  Replaces the
        engine's adapter set."""
        seqs = [s.encode() if isinstance(s, str) else bytes(s) for s in sequences]
        off = np.zeros(len(seqs) + 1, dtype=np.int64)
        if seqs:
            off[1:] = np.cumsum([len(s) for s in seqs])
        pairs = np.ascontiguousarray(pairs, dtype=np.int32).reshape(-1, 2)
        pa, pb_ = np.ascontiguousarray(pairs[:, 0]), np.ascontiguousarray(pairs[:, 1])
        flags = np.zeros(len(pairs), dtype=np.int32)
        self._check(w.C_LIB.pcb_check_compatibility(self._h, b''.join(seqs), off.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)), len(seqs),
                                                    _ptr(pa), _ptr(pb_), len(pairs), _ptr(flags)), 'pcb_check_compatibility')
        self.adapters = seqs
        return flags

    # -- timing / counters ---------------------------------------------------------------------
    def timer_start(self):
        self._check(w.C_LIB.pcb_timer_start(self._h), 'pcb_timer_start')

    def timer_stop(self):
        ms = c_float()
        self._check(w.C_LIB.pcb_timer_stop(self._h, byref(ms)), 'pcb_timer_stop')
        return ms.value

    # -- submit / collect (pcb_set_deferred, pcb_wait) ----------------------------------------------------
    def set_deferred(self, on):
        self._check(w.C_LIB.pcb_set_deferred(self._h, 1 if on else 0), 'pcb_set_deferred')
        self._deferred = bool(on)
        self._held = []

    def wait(self):
        """Completes every queued call; the arrays the align_* calls returned are valid after it."""
        try:
            self._check(w.C_LIB.pcb_wait(self._h), 'pcb_wait')
        finally:
            self._held = []

    def wait_slot(self, slot):
        """Collects one slot's queued align_windows_hits call; everything queued in other slots keeps running."""
        self._check(w.C_LIB.pcb_wait_slot(self._h, int(slot)), 'pcb_wait_slot')

    @contextlib.contextmanager
    def deferred(self):
        """with eng.deferred(): a = eng.align_windows(...); b = eng.align_windows(...)   # a, b valid after the block"""
        self.set_deferred(True)
        try:
            yield self
            self.wait()
        finally:
            self.set_deferred(False)

    def synchronize(self):
        self._check(w.C_LIB.pcb_synchronize(self._h), 'pcb_synchronize')

    def counters(self):
        c = w.PcbCounters()
        self._check(w.C_LIB.pcb_get_counters(self._h, byref(c)), 'pcb_get_counters')
        return {f: getattr(c, f) for f, _ in w.PcbCounters._fields_}

    def reset_counters(self):
        self._check(w.C_LIB.pcb_reset_counters(self._h), 'pcb_reset_counters')

    def set_kernel_timing(self, on):
        self._check(w.C_LIB.pcb_set_kernel_timing(self._h, 1 if on else 0), 'pcb_set_kernel_timing')


def identities(rec):
    """(aligned_region_percent_identity, full_adapter_percent_identity) as src/alignment.cpp:81,89."""
    with np.errstate(divide='ignore', invalid='ignore'):
        aligned = 100.0 * rec['matches'].astype(np.float64) / rec['aligned_len'].astype(np.float64)
        full = 100.0 * rec['matches'].astype(np.float64) / rec['full_len'].astype(np.float64)
    return aligned, full


def to_align_adapter_tuple(rec):
    """
    One record -> what nanopore_read.align_adapter returns (nanopore_read.py:485-500):
    (full_adapter_percent_identity, aligned_region_percent_identity, read_start, read_end [exclusive]).
    The reference parses the printed string, so the identities go through the same "%f" round trip.
    """
    if int(rec['read_start']) == -1:
        return 0.0, 0.0, -1, 0
    parts = w.format_result(rec).split(',')
    return float(parts[6]), float(parts[5]), int(parts[0]), int(parts[1]) + 1
