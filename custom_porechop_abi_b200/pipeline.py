"""
The reference's main() between argument parsing and exit (porechop_abi.py:66-130), on this repo's pieces: load the reads as
one buffer (read_blob.load_reads), phase A on the check reads, the host logic between the phases (the caller's: it depends on
the adapter database, which is the reference's data), phases B and C, the fork's filter, the writer.

    from porechop_abi.adapters import ADAPTERS                      # the reference's database and its own rewrites
    from porechop_abi import porechop_abi as pc
    def between(matching, check_barcodes):
        matching = pc.fix_up_1d2_sets(matching)
        direction = pc.choose_barcoding_kit(matching, 0, None) if check_barcodes else None
        return pc.add_full_barcode_adapter_sets(matching), direction
    pipeline.trim(path, [a for a in ADAPTERS if '(full sequence)' not in a.name], output='out.fastq', between_phases=between)

Every step is the function the tests pin against the reference's own state and bytes (tests/golden/); this module only
strings them together with the reference's defaults (arg_parser.py).
"""
import numpy as np

from . import porechop_batch as pb
from . import read_output as ro
from . import sharding
from .read_blob import load_reads

DEFAULT_SCORING = (3, -6, -5, -2)


def trim(input_file_or_directory, search_adapters, output=None, out_format='auto', barcode_dir=None, between_phases=None,
         engine=None, verbosity=0, print_dest=None, threads=1, check_reads=10000, scoring_scheme_vals=DEFAULT_SCORING,
         end_size=150, adapter_threshold=90.0, end_threshold=75.0, extra_end_trim=2, min_trim_size=4,
         middle_threshold=90.0, extra_middle_trim_good_side=10, extra_middle_trim_bad_side=100, no_split=False,
         discard_middle=False, min_split_read_size=1000, barcode_threshold=75.0, barcode_diff=5.0,
         require_two_barcodes=False, discard_unassigned=False, untrimmed=False, keep_only_reads_with_adapters=True, stdout=None,
         distributed=False, device=None):
    """
    search_adapters: the adapter sets phase A looks for (objects with name, start_sequence, end_sequence, best_start_score,
    best_end_score, best_start_or_end_score(), is_barcode(), barcode_direction(), get_barcode_name(): the reference's
    Adapter).  between_phases(matching_sets, check_barcodes) -> (matching_sets, forward_or_reverse_barcodes): the
    reference's fix_up_1d2_sets / choose_barcoding_kit / add_full_barcode_adapter_sets (porechop_abi.py:78-88); without
    it the sets go on as phase A found them and barcodes count as 'forward'.
    keep_only_reads_with_adapters: the fork's filter_reads_by_adapter (porechop_abi.py:36-39, :121).
    distributed: under an initialised torch.distributed process group every rank takes a length-balanced shard of the reads
    (sharding.balanced_shards) in a buffer of its own, the best identities of phase A are merged (MAX, as the reference's
    max over its check reads), and rank 0 writes everybody's records in read order (sharding.gather_records): the same
    bytes as one process writes.  `device`: where the small merge tensors live ('cuda' under NCCL).
    -> dict(reads, matching_sets, written = what output_reads reports (None on ranks > 0), read_type)
    """
    scoring = list(scoring_scheme_vals)
    blob, reads, check, check_blob, read_type = load_reads(input_file_or_directory, check_reads)
    dist = sharding._dist() if distributed else None
    world = dist.get_world_size() if dist is not None else 1
    global_index = np.arange(len(reads), dtype=np.int64)
    if world > 1:
        mine = np.sort(np.asarray(sharding.balanced_shards(np.asarray(blob.seq_len), world)[dist.get_rank()], dtype=np.int64))
        calls = [reads[i].albacore_barcode_call for i in mine]
        is_check = np.zeros(len(reads), dtype=bool)
        is_check[[r._i for r in check]] = True
        blob = blob.take(mine).compact()
        reads = blob.make_reads()
        for r, call in zip(reads, calls):
            r.albacore_barcode_call = call
        local_check = np.flatnonzero(is_check[mine])
        check, check_blob = [reads[i] for i in local_check], blob.take(local_check)
        global_index = mine
    matching_sets = pb.find_matching_adapter_sets(check, verbosity, end_size, scoring, print_dest, adapter_threshold, threads,
                                                  search_adapters, engine=engine, read_blob=check_blob)
    if world > 1:                                             # every rank saw a part of the check reads: the best of all parts
        best = sharding.merge_best_identities([a.best_start_score for a in search_adapters] + [a.best_end_score for a in search_adapters], device)
        for k, a in enumerate(search_adapters):
            a.best_start_score, a.best_end_score = float(best[k]), float(best[len(search_adapters) + k])
        matching_sets = [a for a in search_adapters if a.best_start_or_end_score() >= adapter_threshold]
    check_barcodes = barcode_dir is not None
    forward_or_reverse = 'forward' if check_barcodes else None
    if between_phases is not None:
        matching_sets, forward_or_reverse = between_phases(matching_sets, check_barcodes)
    if matching_sets:
        pb.find_adapters_at_read_ends(reads, matching_sets, verbosity, end_size, extra_end_trim, end_threshold, scoring, print_dest,
                                      min_trim_size, threads, check_barcodes, barcode_threshold, barcode_diff, require_two_barcodes,
                                      forward_or_reverse, engine=engine, read_blob=blob)
        if not no_split:
            pb.find_adapters_in_read_middles(reads, matching_sets, verbosity, middle_threshold, extra_middle_trim_good_side,
                                             extra_middle_trim_bad_side, scoring, print_dest, threads, discard_middle,
                                             engine=engine, read_blob=blob)
        kept = ro.filter_reads_by_adapter(reads) if keep_only_reads_with_adapters else reads
    else:
        kept = reads                                          # "No adapters found - output reads are unchanged from input reads"
    if world == 1:
        written = ro.output_reads(kept, out_format, output, read_type, verbosity, discard_middle, min_split_read_size, print_dest,
                                  barcode_dir, input_file_or_directory, untrimmed, threads, discard_unassigned, stdout=stdout)
    else:
        idx, records = [], []
        for r in kept:
            mine_of = ro.records_of([r], out_format, output, read_type, discard_middle, min_split_read_size, barcode_dir,
                                    input_file_or_directory, untrimmed, discard_unassigned)
            records.extend(mine_of)
            idx.extend([int(global_index[r._i])] * len(mine_of))
        merged = sharding.gather_records(idx, records, device)
        written = None if merged is None else ro.write_records(merged, out_format, output, read_type, verbosity, print_dest, barcode_dir,
                                                               input_file_or_directory, stdout=stdout)
    return {'reads': reads, 'kept': kept, 'matching_sets': matching_sets, 'written': written, 'read_type': read_type}
