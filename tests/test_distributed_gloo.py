"""CPU, world_size 2 over gloo: the N > 1 host path (shards, MAX / SUM merges, result gather)."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np
    import torch.distributed as dist
    from custom_porechop_abi_b200 import sharding
    dist.init_process_group('gloo')
    rank, world = dist.get_rank(), dist.get_world_size()
    lengths = np.clip(np.random.default_rng(1).lognormal(8.75, 0.75, 1001), 200, 200000).astype(np.int64)
    shards = sharding.balanced_shards(lengths, world)
    mine = shards[rank]
    # stand-in for per-read results: a function of the global read index, so the merge can be checked exactly
    rec = np.stack([mine * 3 + 1, lengths[mine]], axis=1)
    merged = sharding.gather_results(mine, rec, len(lengths))
    best_local = np.full(6, -1.0); best_local[rank::world] = 50.0 + rank
    best = sharding.merge_best_identities(best_local)
    # structured records (the library's 32-byte results) through the tensor gather
    srec = np.zeros(len(mine), dtype=[('a', '<i4'), ('b', '<i4')]); srec['a'] = mine; srec['b'] = lengths[mine] %% 1000
    smerged = sharding.gather_results(mine, srec, len(lengths))
    # phase A, exact: (matches, full_len) pairs, the larger ratio wins, (0, 0) never does
    pairs = np.zeros(4, dtype=[('matches', '<i4'), ('full_len', '<i4')])
    pairs['matches'] = [20, 0, 23, 11] if rank == 0 else [21, 5, 22, 0]
    pairs['full_len'] = [24, 0, 24, 12] if rank == 0 else [25, 30, 24, 0]
    mb = sharding.merge_best(pairs)
    counts = sharding.merge_counters([len(mine), int(lengths[mine].sum())])
    if rank == 0:
        assert merged is not None and np.array_equal(merged[:, 0], np.arange(len(lengths)) * 3 + 1)
        assert np.array_equal(merged[:, 1], lengths)
        assert counts[0] == len(lengths) and counts[1] == lengths.sum()
        expect = np.full(6, -1.0)
        for r in range(world):
            expect[r::world] = 50.0 + r
        assert np.array_equal(best, expect), (best, expect)
        assert np.array_equal(smerged['a'], np.arange(len(lengths))) and np.array_equal(smerged['b'], lengths %% 1000)
        assert mb.tolist() == [[21, 25], [5, 30], [23, 24], [11, 12]], mb.tolist()
        print('GLOO_OK')
    else:
        assert merged is None
    dist.barrier()
    dist.destroy_process_group()
''') % ROOT


def test_world_size_two_gloo(tmp_path):
    script = tmp_path / 'worker.py'
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR='127.0.0.1')
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', '29533', str(script)],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    assert 'GLOO_OK' in out.stdout


def test_bind_to_gpu_cpus_changes_nothing_without_nvml():
    """Without a GPU / NVML the binding helper leaves the process where it is and says so."""
    import os
    from custom_porechop_abi_b200.sharding import bind_to_gpu_cpus
    before = os.sched_getaffinity(0)
    got = bind_to_gpu_cpus(0)
    assert got is None or set(got) <= before
    if got is None:
        assert os.sched_getaffinity(0) == before
    os.sched_setaffinity(0, before)


PIPELINE_WORKER = textwrap.dedent('''
    import hashlib, os, sys, pathlib
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, 'tests'))
    import torch.distributed as dist
    import test_phases as tp
    from custom_porechop_abi_b200 import pipeline
    dist.init_process_group('gloo')
    rank = dist.get_rank()
    tmp = pathlib.Path(sys.argv[1])
    for idx in (0, 5, 9):
        c, g = tp.cases()[idx], tp.out_golden()['cases'][idx]
        p = c['params']
        d = tmp / ('case%%d' %% idx)
        if rank == 0:
            d.mkdir()
            tp._write_input(c, d)
        dist.barrier()
        src = str(d / 'albacore') if (d / 'albacore').exists() else str(next(d.glob('input.*')))
        def between(matching, check_barcodes):
            assert [a.name for a in matching] == c['phase_a']['matching'], (rank, [a.name for a in matching])
            return [tp.AdapterSet(x) for x in c['matching_used']], c['forward_or_reverse']
        for tag in ('file_auto',) + (('bins',) if c['check_barcodes'] else ()):
            a = g['files'][tag]['args']
            out = d / ('out_' + tag)
            if rank == 0:
                out.mkdir()
            dist.barrier()
            res = pipeline.trim(src, [tp.AdapterSet(x) for x in c['search_adapters']], output=str(out / a['output']) if a['output'] else None,
                                out_format=a['out_format'], barcode_dir=str(out / a['barcode_dir']) if a['barcode_dir'] else None,
                                between_phases=between, engine=tp.ReducedOracleEngine(), scoring_scheme_vals=p['scoring'], end_size=p['end_size'],
                                adapter_threshold=p['adapter_threshold'], end_threshold=p['end_threshold'], extra_end_trim=p['extra_end_trim'],
                                min_trim_size=p['min_trim_size'], middle_threshold=p['middle_threshold'],
                                extra_middle_trim_good_side=p['good_side'], extra_middle_trim_bad_side=p['bad_side'],
                                discard_middle=a['discard_middle'], min_split_read_size=a['min_split'], barcode_threshold=p['barcode_threshold'],
                                barcode_diff=p['barcode_diff'], require_two_barcodes=c['require_two_barcodes'],
                                discard_unassigned=a['discard_unassigned'], untrimmed=a['untrimmed'], distributed=True)
            assert 0 < len(res['reads']) < len(c['reads'])                  # a shard, not everything
            dist.barrier()
            if rank == 0:
                where = str(out / a['barcode_dir']) if a['barcode_dir'] else str(out)
                got = {fn: hashlib.sha256(open(os.path.join(where, fn), 'rb').read()).hexdigest() for fn in sorted(os.listdir(where))}
                assert got == g['files'][tag]['sha256'], (idx, tag, sorted(got), sorted(g['files'][tag]['sha256']))
            else:
                assert res['written'] is None
    if rank == 0:
        print('PIPELINE_GLOO_OK')
    dist.barrier()
    dist.destroy_process_group()
''') % (ROOT, ROOT)


def test_sharded_pipeline_writes_the_same_bytes(tmp_path):
    """pipeline.trim on two ranks (gloo, the oracle as engine): shards of the reads, phase A merged over the ranks, rank 0 writes
    what both produced -- the bytes the reference wrote for its fixtures (file, barcode bins, Albacore directory)."""
    script = tmp_path / 'pipeline_worker.py'
    script.write_text(PIPELINE_WORKER)
    work = tmp_path / 'work'
    work.mkdir()
    env = dict(os.environ, MASTER_ADDR='127.0.0.1')
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', '29534', str(script), str(work)],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    assert 'PIPELINE_GLOO_OK' in out.stdout

