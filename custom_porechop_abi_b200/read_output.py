"""
The caller on the far side of the hot path: what the reference does with the state the three phases leave on its reads
(porechop_abi.py:36-39 filter_reads_by_adapter, :535-660 output_reads; nanopore_read.py:66-156 get_fasta / get_fastq /
get_split_read_parts).  A restatement for read objects of either kind -- the reference-shaped ones of the tests, or
read_blob.BlobRead -- with the reference's function names and argument meaning, byte-identical output
(tests/golden/output_golden.json is made by the reference's own functions).

Differences in how, not in what:
 * get_split_read_parts asks `i in middle_trim_positions` for every base of the read in a Python loop; here the set is
   turned into the sorted cut positions inside the trimmed read once and the parts are slices between them.
 * a read is RNA when its sequence held more U than T when it was loaded (nanopore_read.py:36-41); BlobRead answers that
   from the file's bytes, other objects through their `rna` attribute (absent = False).
"""
import gzip
import os
import sys
from collections import defaultdict

import numpy as np


def add_line_breaks_to_sequence(sequence, line_length):
    """Wrapped to line_length, every line (also the last) ends in a line break (misc.py:327-338)."""
    if not sequence:
        return '\n'
    return ''.join(sequence[p:p + line_length] + '\n' for p in range(0, len(sequence), line_length))


def add_number_to_read_name(read_name, number):
    """read_name_N, the number going in front of the first tab or space of the name (nanopore_read.py:503-509)."""
    for sep in ('\t', ' '):
        if sep in read_name:
            return read_name.replace(sep, '_' + str(number) + sep, 1)
    return read_name + '_' + str(number)


def _trimmed(read):
    """(sequence, qualities) with the start / end adapters cut off (nanopore_read.py:66-83; Python slice semantics for
    trims longer than the read, as there)."""
    seq, quals = read.seq, _quals(read)
    if not read.start_trim_amount and not read.end_trim_amount:
        return seq, quals
    return seq[read.start_trim_amount:len(seq) - read.end_trim_amount], quals[read.start_trim_amount:len(quals) - read.end_trim_amount]


def _quals(read):
    q = getattr(read, 'quals', '')
    n = len(read.seq)
    return q + '+' * (n - len(q)) if len(q) < n else q      # padded as NanoporeRead.__init__ pads them (:44-46)


def get_split_read_parts(read, min_split_read_size):
    """The trimmed read cut at its middle_trim_positions: maximal runs of positions that are not in the set, those shorter
    than min_split_read_size dropped (nanopore_read.py:85-104)."""
    seq, quals = _trimmed(read)
    n = len(seq)
    cut = np.fromiter((p for p in read.middle_trim_positions if 0 <= p < n), dtype=np.int64)
    keep = np.ones(n + 2, dtype=bool)
    keep[0] = keep[n + 1] = False
    keep[cut + 1] = False
    edge = np.flatnonzero(keep[1:] != keep[:-1])             # run starts and ends alternate
    parts = [(seq[a:b], quals[a:b]) for a, b in zip(edge[0::2].tolist(), edge[1::2].tolist())]
    return [x for x in parts if len(x[0]) >= min_split_read_size]


def _rna(read):
    return bool(getattr(read, 'rna', False))


def get_fasta(read, min_split_read_size, discard_middle, untrimmed=False):
    """nanopore_read.py:106-130"""
    if not read.middle_trim_positions:
        seq = read.seq if untrimmed else _trimmed(read)[0]
        if not seq:
            return ''
        if _rna(read):
            seq = seq.replace('T', 'U')
        return ''.join(['>', read.name, '\n', add_line_breaks_to_sequence(seq, 70)])
    if discard_middle:
        return ''
    out = []
    for i, (part, _) in enumerate(get_split_read_parts(read, min_split_read_size)):
        if not part:
            return ''
        seq = add_line_breaks_to_sequence(part, 70)
        if _rna(read):
            seq = seq.replace('T', 'U')
        out.append(''.join(['>', add_number_to_read_name(read.name, i + 1), '\n', seq]))
    return ''.join(out)


def get_fastq(read, min_split_read_size, discard_middle, untrimmed=False):
    """nanopore_read.py:132-156"""
    if not read.middle_trim_positions:
        seq, quals = (read.seq, _quals(read)) if untrimmed else _trimmed(read)
        if not seq:
            return ''
        if _rna(read):
            seq = seq.replace('T', 'U')
        return ''.join(['@', read.name, '\n', seq, '\n+\n', quals, '\n'])
    if discard_middle:
        return ''
    out = []
    for i, (seq, qual) in enumerate(get_split_read_parts(read, min_split_read_size)):
        if not seq:
            return ''
        if _rna(read):
            seq = seq.replace('T', 'U')
        out.append(''.join(['@', add_number_to_read_name(read.name, i + 1), '\n', seq, '\n+\n', qual, '\n']))
    return ''.join(out)


def filter_reads_by_adapter(reads):
    """The fork's filter (porechop_abi.py:36-39): only reads with an adapter alignment at BOTH ends go on to the output."""
    return [read for read in reads if (read.adapters_found() if hasattr(read, 'adapters_found')
                                       else bool(read.start_adapter_alignments) and bool(read.end_adapter_alignments))]


def resolve_format(out_format, output, read_type, barcode_dir, input_filename):
    """--format auto (porechop_abi.py:552-572) -> (format without .gz, gzipped?)"""
    if out_format == 'auto':
        if output is None:
            out_format = read_type.lower()
            if barcode_dir is not None and input_filename.lower().endswith('.gz'):
                out_format += '.gz'
        else:
            low = output.lower()
            out_format = next((f for f in ('fasta.gz', 'fastq.gz', 'fasta', 'fastq') if '.' + f in low), read_type.lower())
    gzipped = out_format.endswith('.gz') and (barcode_dir is not None or output is not None)
    if out_format.endswith('.gz') and gzipped:
        out_format = out_format[:-3]
    return out_format, gzipped


def records_of(reads, out_format, output, read_type, discard_middle, min_split_size, barcode_dir, input_filename, untrimmed,
               discard_unassigned):
    """What output_reads would write for these reads, read by read: (barcode bin or '', text, bases counted for the bin's
    summary); reads that produce nothing are left out.  (Apart from output_reads so that ranks can produce their shares and
    one rank writes them: sharding.gather_records.)"""
    out_format, _ = resolve_format(out_format, output, read_type, barcode_dir, input_filename)
    record = get_fasta if out_format == 'fasta' else get_fastq
    out = []
    for read in reads:
        if barcode_dir is not None:                          # bins: `untrimmed` counts, unassigned reads may be dropped (:588-610)
            name = read.barcode_call
            if discard_unassigned and name == 'none':
                continue
            text = record(read, min_split_size, discard_middle, untrimmed)
            if text:
                out.append((name, text, len(read.seq) if untrimmed else len(_trimmed(read)[0])))
        else:                                                # stdout / one file: trimmed, whatever `untrimmed` says (:641-657)
            text = record(read, min_split_size, discard_middle)
            if text:
                out.append(('', text, 0))
    return out


def write_records(records, out_format, output, read_type, verbosity, print_dest, barcode_dir, input_filename, stdout=None):
    """The writing half of output_reads (porechop_abi.py:581-665): barcode bins, stdout or one file."""
    out_format, gzipped = resolve_format(out_format, output, read_type, barcode_dir, input_filename)
    opener = (lambda path: gzip.open(path, 'wt')) if gzipped else (lambda path: open(path, 'wt'))
    bins = {}
    if barcode_dir is not None:
        os.makedirs(barcode_dir, exist_ok=True)
        files, n_reads, n_bases = {}, defaultdict(int), defaultdict(int)
        for name, text, bases in records:
            if name not in files:
                files[name] = opener(os.path.join(barcode_dir, name + '.' + out_format + ('.gz' if gzipped else '')))
            files[name].write(text)
            n_reads[name] += 1
            n_bases[name] += bases
        for name in sorted(files):
            files[name].close()
            bins[name] = (n_reads[name], n_bases[name])
    elif output is None:
        dest = stdout if stdout is not None else sys.stdout
        for _, text, _ in records:
            dest.write(text)
    else:
        with opener(output) as f:
            for _, text, _ in records:
                f.write(text)
    if verbosity > 0 and print_dest is not None:
        print('%d records written' % len(records), file=print_dest, flush=True)
    return {'records': len(records), 'bins': bins}


def output_reads(reads, out_format, output, read_type, verbosity, discard_middle, min_split_size, print_dest, barcode_dir,
                 input_filename, untrimmed, threads, discard_unassigned, stdout=None):
    """
    porechop_abi.py:535-668, same arguments (threads only matters to the reference's external pigz; compression is done
    with Python's gzip here, so the .gz BYTES differ from pigz / gzip output while the text inside is identical).
    Returns {'records': n written, 'bins': {barcode: (reads, bases)}} for the caller's summary.
    """
    records = records_of(reads, out_format, output, read_type, discard_middle, min_split_size, barcode_dir, input_filename,
                         untrimmed, discard_unassigned)
    return write_records(records, out_format, output, read_type, verbosity, print_dest, barcode_dir, input_filename, stdout=stdout)
