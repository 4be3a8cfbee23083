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


def _parsed_identities(rec):
    """(aligned_region_percent_identity, full_adapter_percent_identity) as nanopore_read.align_adapter sees them
    (nanopore_read.py:485-500): the doubles of src/alignment.cpp:81,89 printed with %f and parsed back."""
    m = rec['matches'].astype(np.int64)
    out = []
    for length in (rec['aligned_len'].astype(np.int64), rec['full_len'].astype(np.int64)):
        key = (m << 32) | length
        uniq, inv = np.unique(key, return_inverse=True)
        table = np.empty(len(uniq), dtype=np.float64)
        for k, u in enumerate(uniq):
            mm, ll = int(u >> 32), int(u & 0xFFFFFFFF)
            table[k] = float('%f' % (100.0 * mm / ll)) if ll else float('nan')     # the C++ prints "-nan" for 0.0 / 0
        out.append(table[inv].reshape(key.shape))
    failed = rec['read_start'] == -1
    out[0][failed] = 0.0
    out[1][failed] = 0.0
    return out[0], out[1]


def _read_ends(rec):
    """read_end as align_adapter returns it: exclusive, 0 for the failed record."""
    return np.where(rec['read_start'] == -1, 0, rec['read_end'] + 1)


def _windows(reads, end_size, at_end):
    """The reference's seq[:end_size] / seq[-end_size:] (nanopore_read.py:164,170,181,203) as a fixed-stride buffer."""
    stride = max(16, -(-end_size // 16) * 16)
    buf = np.zeros((len(reads), stride), dtype=np.uint8)
    lens = np.empty(len(reads), dtype=np.int32)
    for i, r in enumerate(reads):
        w = (r.seq[-end_size:] if at_end else r.seq[:end_size]).encode()
        lens[i] = len(w)
        buf[i, :len(w)] = np.frombuffer(w, dtype=np.uint8)
    return buf, lens


def _collect_after(eng):
    """Submit / collect around a group of engine calls (AlignmentEngine.deferred); a plain engine collects per call."""
    return eng.deferred() if hasattr(eng, 'deferred') else contextlib.nullcontext()


def _engine(engine, scoring_scheme_vals):
    eng = engine if engine is not None else AlignmentEngine(0)
    eng.set_scoring(scoring_scheme_vals)
    return eng


def find_matching_adapter_sets(check_reads, verbosity, end_size, scoring_scheme_vals, print_dest, adapter_threshold,
                               threads, search_adapters, engine=None):
    """
    Phase A (porechop_abi.py:200-245): every check read's start / end window against every adapter set's start / end
    sequence; each set keeps the best full-adapter identity seen (nanopore_read.py:158-173).  Returns the sets whose
    best_start_or_end_score() reaches adapter_threshold, in search order, as the reference does.
    """
    eng = _engine(engine, scoring_scheme_vals)
    with_start = [a for a in search_adapters if a.start_sequence]
    with_end = [a for a in search_adapters if a.end_sequence]
    seqs = [a.start_sequence[1] for a in with_start] + [a.end_sequence[1] for a in with_end]
    if check_reads and seqs:
        eng.set_adapters(seqs)
        submitted = {}
        with _collect_after(eng):                          # both submits queued, then collected together
            for slot, (at_end, sets, id0) in enumerate(((False, with_start, 0), (True, with_end, len(with_start)))):
                if not sets:
                    continue
                if hasattr(eng, 'select_slot'):
                    eng.select_slot(slot)
                buf, lens = _windows(check_reads, end_size, at_end)
                submitted[at_end] = eng.align_windows(buf, lens, np.arange(id0, id0 + len(sets), dtype=np.int32))
        if hasattr(eng, 'select_slot'):
            eng.select_slot(0)
        for at_end, sets in ((False, with_start), (True, with_end)):
            if at_end not in submitted:
                continue
            rec = submitted[at_end]
            _, full = _parsed_identities(rec)
            best = full.max(axis=0)                       # no NaN here: full_len >= 1 whenever the alignment exists
            for a, score in zip(sets, best):
                if at_end:
                    a.best_end_score = max(a.best_end_score, float(score))
                else:
                    a.best_start_score = max(a.best_start_score, float(score))
    return [x for x in search_adapters if x.best_start_or_end_score() >= adapter_threshold]


def find_adapters_at_read_ends(reads, matching_sets, verbosity, end_size, extra_trim_size, end_threshold,
                               scoring_scheme_vals, print_dest, min_trim_size, threads, check_barcodes,
                               barcode_threshold, barcode_diff, require_two_barcodes, forward_or_reverse_barcodes,
                               engine=None):
    """
    Phase B (porechop_abi.py:359-438): find_start_trim + find_end_trim (nanopore_read.py:175-217) for every read, then
    determine_barcode.  Start windows x start sequences and end windows x end sequences are two GPU submits.
    """
    eng = _engine(engine, scoring_scheme_vals)
    starts = [a for a in matching_sets if a.start_sequence]
    ends = [a for a in matching_sets if a.end_sequence]
    seqs = [a.start_sequence[1] for a in starts] + [a.end_sequence[1] for a in ends]
    if reads and seqs:
        eng.set_adapters(seqs)
    # both submits are queued before either is collected (submit / collect: the copies of the second run under the
    # kernels of the first), as the reference queues all reads on its pool before it collects (porechop_abi.py:395-407)
    submitted = {}
    with _collect_after(eng):
        for slot, (at_end, sets, id0) in enumerate(((False, starts, 0), (True, ends, len(starts)))):
            if not reads or not sets:
                continue
            if hasattr(eng, 'select_slot'):
                eng.select_slot(slot)
            buf, lens = _windows(reads, end_size, at_end)
            submitted[at_end] = eng.align_windows(buf, lens, np.arange(id0, id0 + len(sets), dtype=np.int32))
    if hasattr(eng, 'select_slot'):
        eng.select_slot(0)
    for at_end, sets, id0 in ((False, starts, 0), (True, ends, len(starts))):
        if at_end not in submitted:
            continue
        rec = submitted[at_end]
        partial, full = _parsed_identities(rec)
        r_start, r_end = rec['read_start'], _read_ends(rec)
        with np.errstate(invalid='ignore'):
            edge_ok = (r_start != 0) if at_end else (r_end != end_size)
            hit = (partial > end_threshold) & edge_ok & (r_end - r_start >= min_trim_size)
        for i, k in zip(*np.nonzero(hit)):                 # row-major: per read, in adapter order, as the reference appends
            read, adapter = reads[i], sets[k]
            entry = (adapter, float(full[i, k]), float(partial[i, k]), int(r_start[i, k]), int(r_end[i, k]))
            if at_end:
                read.end_trim_amount = max(read.end_trim_amount, (end_size - int(r_start[i, k])) + extra_trim_size)
                read.end_adapter_alignments.append(entry)
            else:
                read.start_trim_amount = max(read.start_trim_amount, int(r_end[i, k]) + extra_trim_size)
                read.start_adapter_alignments.append(entry)
        if check_barcodes:
            for k, adapter in enumerate(sets):
                if adapter.is_barcode() and adapter.barcode_direction() == forward_or_reverse_barcodes:
                    name = adapter.get_barcode_name()
                    for i, read in enumerate(reads):
                        (read.end_barcode_scores if at_end else read.start_barcode_scores)[name] = float(full[i, k])
    if check_barcodes:
        for read in reads:
            read.determine_barcode(barcode_threshold, barcode_diff, require_two_barcodes)


def find_adapters_in_read_middles(reads, matching_sets, verbosity, middle_threshold, extra_trim_good_side,
                                  extra_trim_bad_side, scoring_scheme_vals, print_dest, threads, discard_middle,
                                  engine=None):
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
    eng.set_adapters([seq for _, seq in adapters])
    n_ad = len(adapters)

    masked = [bytearray(r.get_seq_with_start_end_adapters_trimmed().encode()) for r in reads]
    queue = [(i, 0) for i in range(len(reads))]             # (read, first adapter still to be aligned)
    while queue:
        # one submit: every queued read against its remaining adapters
        lens = np.array([len(masked[i]) for i, _ in queue], dtype=np.int64)
        offs = np.zeros(len(queue) + 1, dtype=np.int64)
        np.cumsum(lens, out=offs[1:])
        blob = np.empty(int(offs[-1]), dtype=np.uint8)
        seg_off, seg_len, seg_ad, first = [], [], [], []
        for q, (i, k0) in enumerate(queue):
            blob[offs[q]:offs[q + 1]] = np.frombuffer(bytes(masked[i]), dtype=np.uint8)
            first.append(len(seg_ad))
            for k in range(k0, n_ad):
                seg_off.append(offs[q]); seg_len.append(lens[q]); seg_ad.append(k)
        rec = eng.align_tasks(blob, np.array(seg_off, dtype=np.int64), np.array(seg_len, dtype=np.int32),
                              np.array(seg_ad, dtype=np.int32))
        _, full = _parsed_identities(rec)
        r_start, r_end = rec['read_start'], _read_ends(rec)
        requeue = []
        for q, (i, k0) in enumerate(queue):
            t0 = first[q]
            for k in range(k0, n_ad):
                t = t0 + (k - k0)
                if full[t] >= middle_threshold:             # a hit: record, mask, align the same adapter again
                    read = reads[i]
                    adapter_name = adapters[k][0]
                    read_start, read_end = int(r_start[t]), int(r_end[t])
                    masked[i][read_start:read_end] = b'-' * (read_end - read_start)
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
                    requeue.append((i, k))
                    break                                   # later adapters were aligned on the old mask: redo them
        queue = requeue
