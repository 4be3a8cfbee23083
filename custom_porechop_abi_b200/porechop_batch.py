"""
Batched replacements for the three per-read loops of the reference that call adapter_alignment
(porechop_abi/porechop_abi.py):

    find_matching_adapter_sets      :200-245   phase A  -> NanoporeRead.align_adapter_set      nanopore_read.py:158-173
    find_adapters_at_read_ends      :359-438   phase B  -> find_start_trim / find_end_trim      nanopore_read.py:175-217
    find_adapters_in_read_middles   :457-522   phase C  -> find_middle_adapters                 nanopore_read.py:219-252

Same names, same arguments (plus the engine, and for phase A the list of adapter sets to search, which the reference
takes from its module-level database), same effects on the read / adapter objects: the functions work on the
reference's own NanoporeRead and Adapter objects (duck-typed: only the attributes those methods touch are used), so a
maintainer swaps three functions and keeps everything else.  What changes is the order of work, not the results: every
(window, adapter) pair of a phase is one GPU submit (AlignmentEngine), and the per-read decisions are then applied in
the reference's order.

Exactness notes
* The reference parses the C string, so its identities are float("%f" % value); `_parsed_identities` reproduces that
  round trip per distinct (matches, length) pair.
* Phase C masks a hit with '-' and aligns the SAME adapter again before moving on (nanopore_read.py:229-252).  Batched
  form: round 0 aligns every (read, adapter) on the unmasked sequence; a read keeps the results of the adapters before
  its first hit (its mask was still empty there), applies the hit, and is re-queued from that adapter on; rounds repeat
  until no read has a hit.  Each result a read consumes was computed on exactly the sequence the reference would
  have passed at that point of its loop.
* Verbosity-dependent printing is the reference's business and is left to the caller (the alignments are stored on the
  reads exactly as the reference stores them, so its formatting methods work unchanged).
"""
import contextlib

import numpy as np

from .engine import AlignmentEngine


def _percent_table(matches, length):
    """float('%f' % (100.0 * m / l)) for every element, through a table over the distinct (m, l) pairs: the C++ side prints
    the double with std::to_string and the reference parses it back (alignment.cpp:112-120, nanopore_read.py:487-498), so
    the identities the thresholds see carry six decimals.  0 / 0 prints "-nan" there: NaN here."""
    m = matches.astype(np.int64)
    length = length.astype(np.int64)
    if m.size == 0:
        return np.zeros(m.shape, dtype=np.float64)
    width = int(length.max()) + 1
    key = m * width + length
    n_keys = int(key.max()) + 1
    if n_keys <= (1 << 22):                                 # the usual case: a few thousand distinct pairs in a small range
        present = np.zeros(n_keys, dtype=bool)
        present[key.ravel()] = True
        table = np.zeros(n_keys, dtype=np.float64)
        for u in np.nonzero(present)[0]:
            mm, ll = divmod(int(u), width)
            table[u] = float('%f' % (100.0 * mm / ll)) if ll else float('nan')
        return table[key]
    uniq, inv = np.unique(key, return_inverse=True)
    table = np.empty(len(uniq), dtype=np.float64)
    for k, u in enumerate(uniq):
        mm, ll = divmod(int(u), width)
        table[k] = float('%f' % (100.0 * mm / ll)) if ll else float('nan')
    return table[inv].reshape(key.shape)


def _parsed_identities(rec, aligned=True):
    """(aligned_region_percent_identity, full_adapter_percent_identity) as nanopore_read.align_adapter sees them
    (nanopore_read.py:485-500): the doubles of src/alignment.cpp:81,89 printed with %f and parsed back.
    aligned=False skips the first one (phases A and C only look at the full-adapter identity)."""
    failed = rec['read_start'] == -1
    full = _percent_table(rec['matches'], rec['full_len'])
    full[failed] = 0.0
    part = None
    if aligned:
        part = _percent_table(rec['matches'], rec['aligned_len'])
        part[failed] = 0.0
    return part, full


def _read_ends(rec):
    """read_end as align_adapter returns it: exclusive, 0 for the failed record."""
    return np.where(rec['read_start'] == -1, 0, rec['read_end'] + 1)


def _windows(reads, end_size, at_end):
    """The reference's seq[:end_size] / seq[-end_size:] (nanopore_read.py:164,170,181,203) as a fixed-stride buffer."""
    stride = max(16, -(-end_size // 16) * 16)
    parts = [r.seq[-end_size:] for r in reads] if at_end else [r.seq[:end_size] for r in reads]
    lens = np.fromiter((len(w) for w in parts), dtype=np.int32, count=len(parts))
    buf = np.zeros((len(reads), stride), dtype=np.uint8)
    if len(parts):
        flat = np.frombuffer(''.join(parts).encode(), dtype=np.uint8)     # windows back to back, row-major like the mask
        buf[np.arange(stride, dtype=np.int32)[None, :] < lens[:, None]] = flat
    return buf, lens


def _trimmed_bounds(r):
    """[lo, hi) of get_seq_with_start_end_adapters_trimmed() inside r.seq (nanopore_read.py:77-81), Python slice rules included."""
    if not r.start_trim_amount and not r.end_trim_amount:
        return 0, len(r.seq)
    lo, hi, _ = slice(r.start_trim_amount, len(r.seq) - r.end_trim_amount).indices(len(r.seq))
    return lo, max(lo, hi)


def _collect_after(eng):
    """Submit / collect around a group of engine calls (AlignmentEngine.deferred); a plain engine collects per call."""
    return eng.deferred() if hasattr(eng, 'deferred') else contextlib.nullcontext()


def _stage_windows(eng, reads, read_blob, end_size, at_end):
    """The windows of a phase for the engine: (buffer, lengths) packed on the host, or (None, n) when the engine can cut
    them out of the read blob on the device (the blob is announced once and serves all three phases)."""
    if read_blob is not None and hasattr(eng, 'stage_windows_from_reads'):
        # (the engine holds the announced array itself: identity of a live object, not an id() that a freed one may pass on)
        if getattr(eng, '_uploaded', None) is not read_blob.buf:
            eng.upload_reads(read_blob.buf)
        return None, eng.stage_windows_from_reads(read_blob.seq_off, read_blob.seq_len, end_size, at_end)
    return read_blob.windows(end_size, at_end) if read_blob is not None else _windows(reads, end_size, at_end)


def _engine(engine, scoring_scheme_vals):
    eng = engine if engine is not None else AlignmentEngine(0)
    eng.set_scoring(scoring_scheme_vals)
    return eng


def _ratio_percent(matches, length):
    """float('%f' % (100.0 * m / l)) per element (see _percent_table); 0.0 where l == 0 (no alignment)."""
    out = _percent_table(matches, np.maximum(length, 1))
    out[np.asarray(length) == 0] = 0.0
    return out


def find_matching_adapter_sets(check_reads, verbosity, end_size, scoring_scheme_vals, print_dest, adapter_threshold,
                               threads, search_adapters, engine=None, read_blob=None):
    """
    Phase A (porechop_abi.py:200-245): every check read's start / end window against every adapter set's start / end
    sequence; each set keeps the best full-adapter identity seen (nanopore_read.py:158-173).  Returns the sets whose
    best_start_or_end_score() reaches adapter_threshold, in search order, as the reference does.
    The maximum is taken on the device (pcb_align_windows_best): 8 bytes per sequence come back, not 32 per pair.
    """
    eng = _engine(engine, scoring_scheme_vals)
    with_start = [a for a in search_adapters if a.start_sequence]
    with_end = [a for a in search_adapters if a.end_sequence]
    seqs = [a.start_sequence[1] for a in with_start] + [a.end_sequence[1] for a in with_end]
    if check_reads and seqs:
        eng.set_adapters(seqs)
        reduced = hasattr(eng, 'align_windows_best')
        submitted = {}
        with _collect_after(eng):                          # both submits queued, then collected together
            for slot, (at_end, sets, id0) in enumerate(((False, with_start, 0), (True, with_end, len(with_start)))):
                if not sets:
                    continue
                if hasattr(eng, 'select_slot'):
                    eng.select_slot(slot)
                ids = np.arange(id0, id0 + len(sets), dtype=np.int32)
                if reduced:
                    buf, lens = _stage_windows(eng, check_reads, read_blob, end_size, at_end)
                    submitted[at_end] = eng.align_windows_best(buf, lens, ids)
                else:
                    buf, lens = read_blob.windows(end_size, at_end) if read_blob is not None else _windows(check_reads, end_size, at_end)
                    submitted[at_end] = eng.align_windows(buf, lens, ids)
        if hasattr(eng, 'select_slot'):
            eng.select_slot(0)
        for at_end, sets in ((False, with_start), (True, with_end)):
            if at_end not in submitted:
                continue
            rec = submitted[at_end]
            if reduced:
                best = _ratio_percent(rec['matches'], rec['full_len'])
            else:
                _, full = _parsed_identities(rec, aligned=False)
                best = full.max(axis=0)                   # no NaN here: full_len >= 1 whenever the alignment exists
            for a, score in zip(sets, best):
                if at_end:
                    a.best_end_score = max(a.best_end_score, float(score))
                else:
                    a.best_start_score = max(a.best_start_score, float(score))
    return [x for x in search_adapters if x.best_start_or_end_score() >= adapter_threshold]


def determine_barcode(read, barcode_threshold, barcode_diff, require_two_barcodes):
    """NanoporeRead.determine_barcode (nanopore_read.py:408-482) for read classes that do not bring their own: best and
    second-best barcode per end (Python's stable descending sort: the first in insertion order among equals), then the
    reference's rules for --require_two_barcodes or for the combined list, then the agreement with Albacore's call."""
    start = sorted(read.start_barcode_scores.items(), reverse=True, key=lambda x: x[1])
    end = sorted(read.end_barcode_scores.items(), reverse=True, key=lambda x: x[1])
    none = ('none', 0.0)
    read.best_start_barcode = start[0] if len(start) >= 1 else getattr(read, 'best_start_barcode', none)
    read.second_best_start_barcode = start[1] if len(start) >= 2 else getattr(read, 'second_best_start_barcode', none)
    read.best_end_barcode = end[0] if len(end) >= 1 else getattr(read, 'best_end_barcode', none)
    read.second_best_end_barcode = end[1] if len(end) >= 2 else getattr(read, 'second_best_end_barcode', none)
    if require_two_barcodes:
        ok = (read.best_start_barcode[1] >= barcode_threshold and read.best_end_barcode[1] >= barcode_threshold and
              read.best_start_barcode[1] >= read.second_best_start_barcode[1] + barcode_diff and
              read.best_end_barcode[1] >= read.second_best_end_barcode[1] + barcode_diff and
              read.best_start_barcode[0] == read.best_end_barcode[0])
        read.barcode_call = read.best_start_barcode[0] if ok else 'none'
    else:
        combined, seen = [], set()
        for name, score in sorted(start + end, reverse=True, key=lambda x: x[1]):
            if name not in seen:
                combined.append((name, score)); seen.add(name)
        best = combined[0] if len(combined) >= 1 else none
        second = combined[1] if len(combined) >= 2 else none
        read.barcode_call = best[0] if (best[1] >= barcode_threshold and best[1] >= second[1] + barcode_diff) else 'none'
    if getattr(read, 'albacore_barcode_call', None) is not None and read.barcode_call != read.albacore_barcode_call:
        read.barcode_call = 'none'


def find_adapters_at_read_ends(reads, matching_sets, verbosity, end_size, extra_trim_size, end_threshold,
                               scoring_scheme_vals, print_dest, min_trim_size, threads, check_barcodes,
                               barcode_threshold, barcode_diff, require_two_barcodes, forward_or_reverse_barcodes,
                               engine=None, read_blob=None):
    """
    Phase B (porechop_abi.py:359-438): find_start_trim + find_end_trim (nanopore_read.py:175-217) for every read, then
    determine_barcode.  Start windows x start sequences and end windows x end sequences are two GPU submits; what comes
    back is already reduced on the device (pcb_align_windows_hits): the pairs the two methods keep, in the reference's
    order, and -- when barcodes are checked -- one 16-bit identity per (read, barcode sequence).
    read_blob: the same reads as a read_blob.ReadBlob (same order); the windows then come out of its buffer with one
    numpy gather instead of a Python slice per read.
    """
    eng = _engine(engine, scoring_scheme_vals)
    starts = [a for a in matching_sets if a.start_sequence]
    ends = [a for a in matching_sets if a.end_sequence]
    seqs = [a.start_sequence[1] for a in starts] + [a.end_sequence[1] for a in ends]
    if reads and seqs:
        eng.set_adapters(seqs)
    reduced = hasattr(eng, 'align_windows_hits')

    def wanted(sets):                                      # the sets whose full identity every read records (nanopore_read.py:190-195)
        return np.array([1 if (check_barcodes and a.is_barcode() and a.barcode_direction() == forward_or_reverse_barcodes) else 0
                         for a in sets], dtype=np.uint8)
    # both submits are queued before either is collected (submit / collect: the copies of the second run under the
    # kernels of the first), as the reference queues all reads on its pool before it collects (porechop_abi.py:395-407)
    submitted = {}
    with _collect_after(eng):
        for slot, (at_end, sets, id0) in enumerate(((False, starts, 0), (True, ends, len(starts)))):
            if not reads or not sets:
                continue
            if hasattr(eng, 'select_slot'):
                eng.select_slot(slot)
            ids = np.arange(id0, id0 + len(sets), dtype=np.int32)
            if reduced:                                    # room for every pair when the list is short, else 16 per read (and once more if not)
                buf, lens = _stage_windows(eng, reads, read_blob, end_size, at_end)
                submitted[at_end] = eng.align_windows_hits(buf, lens, ids, end_threshold, at_end, end_size, min_trim_size,
                                                           want_ident=wanted(sets), hits_cap=len(reads) * min(len(sets), 16))
            else:
                buf, lens = read_blob.windows(end_size, at_end) if read_blob is not None else _windows(reads, end_size, at_end)
                submitted[at_end] = eng.align_windows(buf, lens, ids)
    if hasattr(eng, 'select_slot'):
        eng.select_slot(0)
    for at_end, sets, id0 in ((False, starts, 0), (True, ends, len(starts))):
        if at_end not in submitted:
            continue
        want = wanted(sets)
        if reduced:
            res = submitted[at_end]
            if res.n_hits > len(res.hits):                 # more hits than the deferred call had room for: once more, alone
                buf, lens = _stage_windows(eng, reads, read_blob, end_size, at_end)
                res = eng.align_windows_hits(buf, lens, np.arange(id0, id0 + len(sets), dtype=np.int32), end_threshold, at_end,
                                             end_size, min_trim_size, want_ident=want, hits_cap=res.n_hits)
            h = res.hits
            partial = _ratio_percent(h['matches'], h['aligned_len'])
            full = _ratio_percent(h['matches'], h['full_len'])
            ok = partial > end_threshold                   # the device keeps a tie at six decimals; the parsed value decides
            ii, kk = h['window'][ok], h['id_pos'][ok]
            fulls, parts = full[ok], partial[ok]
            starts_, ends_ = h['read_start'][ok], h['read_end'][ok] + 1
            if check_barcodes and want.any():
                m, length = res.identities()
                scores = _ratio_percent(m, length)         # [reads, wanted sets]
        else:
            rec = submitted[at_end]
            partial, full = _parsed_identities(rec)
            r_start, r_end = rec['read_start'], _read_ends(rec)
            with np.errstate(invalid='ignore'):
                edge_ok = (r_start != 0) if at_end else (r_end != end_size)
                hit = (partial > end_threshold) & edge_ok & (r_end - r_start >= min_trim_size)
            ii, kk = np.nonzero(hit)                       # row-major: per read, in adapter order, as the reference appends
            fulls, parts = full[ii, kk], partial[ii, kk]
            starts_, ends_ = r_start[ii, kk], r_end[ii, kk]
            if check_barcodes and want.any():
                scores = full[:, np.flatnonzero(want)]
        if read_blob is not None and read_blob.owns(reads):
            # the reads are views of the blob: trims and alignment lists for all of them at once
            read_blob.apply_end_hits(at_end, sets, np.asarray(ii), np.asarray(kk), np.asarray(fulls), np.asarray(parts),
                                     np.asarray(starts_, dtype=np.int64), np.asarray(ends_, dtype=np.int64), end_size, extra_trim_size)
            ii = kk = fulls = parts = starts_ = ends_ = np.zeros(0, dtype=np.int64)
        for i, k, f, pa, rs_, re_ in zip(ii.tolist(), kk.tolist(), np.asarray(fulls).tolist(), np.asarray(parts).tolist(),
                                         np.asarray(starts_).tolist(), np.asarray(ends_).tolist()):
            read = reads[i]
            entry = (sets[k], f, pa, rs_, re_)
            if at_end:
                read.end_trim_amount = max(read.end_trim_amount, (end_size - rs_) + extra_trim_size)
                read.end_adapter_alignments.append(entry)
            else:
                read.start_trim_amount = max(read.start_trim_amount, re_ + extra_trim_size)
                read.start_adapter_alignments.append(entry)
        if check_barcodes and want.any():
            # one dict per read, keys in adapter order (a later set with the same barcode name overwrites the value, not
            # the position, as the reference's assignment does)
            names = [a.get_barcode_name() for a, wt in zip(sets, want) if wt]
            rows = scores.tolist()
            attr = 'end_barcode_scores' if at_end else 'start_barcode_scores'
            for read, row in zip(reads, rows):
                getattr(read, attr).update(zip(names, row))
    if check_barcodes:
        for read in reads:
            if hasattr(read, 'determine_barcode'):
                read.determine_barcode(barcode_threshold, barcode_diff, require_two_barcodes)
            else:
                determine_barcode(read, barcode_threshold, barcode_diff, require_two_barcodes)


def find_adapters_in_read_middles(reads, matching_sets, verbosity, middle_threshold, extra_trim_good_side,
                                  extra_trim_bad_side, scoring_scheme_vals, print_dest, threads, discard_middle,
                                  engine=None, read_blob=None):
    """
    Phase C (porechop_abi.py:457-522): NanoporeRead.find_middle_adapters (nanopore_read.py:219-252) for every read, in
    rounds (see the module docstring).  Updates middle_adapter_positions, middle_trim_positions and middle_hit_str.
    """
    eng = _engine(engine, scoring_scheme_vals)
    adapters = []                                           # (name, sequence) in the reference's order (:463-470)
    for matching_set in matching_sets:
        if matching_set.start_sequence:
            adapters.append(matching_set.start_sequence)
        if matching_set.end_sequence:
            if (not matching_set.start_sequence) or matching_set.end_sequence[1] != matching_set.start_sequence[1]:
                adapters.append(matching_set.end_sequence)
    start_sequence_names = set(m.start_sequence[0] for m in matching_sets if m.start_sequence)
    end_sequence_names = set(m.end_sequence[0] for m in matching_sets if m.end_sequence)
    if not reads or not adapters:
        return
    if read_blob is not None and len(read_blob) != len(reads):
        raise ValueError('read_blob holds %d reads, the read list %d' % (len(read_blob), len(reads)))
    eng.set_adapters([seq for _, seq in adapters])
    n_ad = len(adapters)

    # A read's sequence as the reference's loop sees it: get_seq_with_start_end_adapters_trimmed() (nanopore_read.py:77-81),
    # with every hit so far overwritten by '-'.  Round 0 does not slice a billion bases in Python: the untrimmed reads go
    # up as one blob and the trimmed sequence is a segment of it, with Python's own slice arithmetic (a trim longer than
    # the read makes the reference's stop index negative, which Python counts from the end).  A read becomes a bytearray
    # of its trimmed sequence once it has a hit.
    current = [None] * len(reads)                           # bytearray of the masked trimmed sequence, once a read has a hit
    q_read = np.arange(len(reads), dtype=np.int64)          # queued reads and, per read, the first adapter still to align
    q_k0 = np.zeros(len(reads), dtype=np.int64)
    first_round = True
    while len(q_read):
        # one submit: every queued read against its remaining adapters
        if first_round and read_blob is not None:           # the file's own bytes are the blob: nothing to join or encode
            blob = read_blob.buf
            if read_blob.owns(reads):
                offs, lens = read_blob.trimmed_segments(read_blob.start_trim, read_blob.end_trim)
            else:
                offs, lens = read_blob.trimmed_segments(
                    np.fromiter((r.start_trim_amount for r in reads), dtype=np.int64, count=len(reads)),
                    np.fromiter((r.end_trim_amount for r in reads), dtype=np.int64, count=len(reads)))
            first_round = False
        elif first_round:
            bounds = np.array([_trimmed_bounds(r) for r in reads], dtype=np.int64).reshape(len(reads), 2)
            full_lens = np.fromiter((len(r.seq) for r in reads), dtype=np.int64, count=len(reads))
            base = np.zeros(len(reads) + 1, dtype=np.int64)
            np.cumsum(full_lens, out=base[1:])
            blob = np.frombuffer(''.join(r.seq for r in reads).encode(), dtype=np.uint8)
            offs = base[:-1] + bounds[:, 0]
            lens = bounds[:, 1] - bounds[:, 0]
            first_round = False
        else:
            parts = [bytes(current[i]) for i in q_read]
            lens = np.fromiter((len(x) for x in parts), dtype=np.int64, count=len(parts))
            offs = np.zeros(len(parts) + 1, dtype=np.int64)
            np.cumsum(lens, out=offs[1:])
            offs = offs[:-1]
            blob = np.frombuffer(b''.join(parts), dtype=np.uint8)
        counts = n_ad - q_k0                                 # tasks per queued read
        first = np.zeros(len(q_read) + 1, dtype=np.int64)
        np.cumsum(counts, out=first[1:])
        entry = np.repeat(np.arange(len(q_read), dtype=np.int64), counts)    # queue entry of every task
        seg_ad = np.arange(int(first[-1]), dtype=np.int64) - first[entry] + q_k0[entry]
        # (the reads may be on the device already: phases A / B announced the same blob for their windows)
        resident = read_blob is not None and blob is read_blob.buf and getattr(eng, '_uploaded', None) is blob
        rec = eng.align_tasks(None if resident else blob, offs[entry], lens[entry].astype(np.int32), seg_ad.astype(np.int32))
        _, full = _parsed_identities(rec, aligned=False)
        r_start, r_end = rec['read_start'], _read_ends(rec)
        # first hit of every queued read (tasks are in adapter order per read); NaN compares False, as in the reference
        with np.errstate(invalid='ignore'):
            hits = np.nonzero(full >= middle_threshold)[0]
        hit_entry, pos = np.unique(entry[hits], return_index=True)
        next_read, next_k0 = [], []
        for q, t in zip(hit_entry, hits[pos]):               # a hit: record, mask, align the same adapter again
            i, k = int(q_read[q]), int(seg_ad[t])
            read = reads[i]
            adapter_name = adapters[k][0]
            read_start, read_end = int(r_start[t]), int(r_end[t])
            if current[i] is None:
                current[i] = bytearray(read.get_seq_with_start_end_adapters_trimmed().encode())
            current[i][read_start:read_end] = b'-' * (read_end - read_start)
            read.middle_adapter_positions.update(range(read_start, read_end))
            read.middle_hit_str += '  ' + adapter_name + ' (read coords: ' + str(read_start) + '-' + \
                str(read_end) + ', ' + 'identity: ' + '%.1f' % full[t] + '%)\n'
            trim_start = read_start - extra_trim_good_side
            if adapter_name in start_sequence_names:
                trim_start = read_start - extra_trim_bad_side
            trim_end = read_end + extra_trim_good_side
            if adapter_name in end_sequence_names:
                trim_end = read_end + extra_trim_bad_side
            read.middle_trim_positions.update(range(trim_start, trim_end))
            next_read.append(i); next_k0.append(k)           # later adapters were aligned on the old mask: redo them
        q_read = np.array(next_read, dtype=np.int64)
        q_k0 = np.array(next_k0, dtype=np.int64)
