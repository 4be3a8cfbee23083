"""GPU, world_size 2 over NCCL: ONE read set dealt to two ranks by sharding.balanced_shards reproduces, after the NCCL
merges (phase-A best identities, summary counters) and the tensor gather of the per-read records, exactly what one GPU
computes on the whole set.  Needs two devices (the round-end GPU tier has one: run with `gpurun --gpus 2`, log under
profiles/)."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, %r)
    import numpy as np
    import torch
    import torch.distributed as dist
    from custom_porechop_abi_b200 import sharding, synthetic
    from custom_porechop_abi_b200.engine import AlignmentEngine
    local = int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    rank, world = dist.get_rank(), dist.get_world_size()
    N = 20000
    adapters, s_ids, e_ids = synthetic.adapter_set()
    lengths = synthetic.read_lengths(N, np.random.default_rng(3), hi=30000)
    blob, off = synthetic.make_reads(N, adapters, s_ids, e_ids, seed=4, lengths=lengths)      # the whole set, same on every rank
    sbuf, ebuf, wlen = synthetic.end_windows(blob, off, 150, 160)
    ids = np.array(s_ids[:40], dtype=np.int32)
    mid = np.array([s_ids[0], e_ids[0]], dtype=np.int32)
    shards = sharding.balanced_shards(lengths, world)
    loads = np.array([lengths[s].sum() for s in shards], dtype=np.float64)
    assert loads.max() / loads.mean() < 1.02
    mine = shards[rank]

    def run(eng, sel):
        """phase A best, phase B records + hits, phase C records for the reads `sel`"""
        best = eng.align_windows_best(sbuf[sel], wlen[sel], ids)
        rec = eng.align_windows(sbuf[sel], wlen[sel], ids)
        hits = eng.align_windows_hits(sbuf[sel], wlen[sel], ids, 75.0, False, 150, 4)
        seg_off = np.repeat(off[:-1][sel], len(mid)); seg_len = np.repeat(lengths[sel].astype(np.int32), len(mid))
        mrec = eng.align_tasks(blob, seg_off, seg_len, np.tile(mid, len(sel))).reshape(len(sel), len(mid))
        return best, rec, hits.n_hits, mrec

    eng = AlignmentEngine(local)
    eng.set_adapters(adapters)
    best, rec, n_hits, mrec = run(eng, mine)
    merged_best = sharding.merge_best(best, device='cuda')
    counts = sharding.merge_counters([len(mine), n_hits], device='cuda')
    all_rec = sharding.gather_results(mine, rec, N, device='cuda')
    all_mrec = sharding.gather_results(mine, mrec, N, device='cuda')
    if rank == 0:
        best1, rec1, n_hits1, mrec1 = run(eng, np.arange(N))
        assert all_rec.tobytes() == rec1.tobytes() and all_mrec.tobytes() == mrec1.tobytes()
        assert counts.tolist() == [N, n_hits1]
        # the same best identity per adapter (the pair may differ when two pairs have the same ratio)
        a = merged_best[:, 0] * np.maximum(best1['full_len'].astype(np.int64), 1)
        b = best1['matches'].astype(np.int64) * np.maximum(merged_best[:, 1], 1)
        assert np.array_equal(a, b), (merged_best[:5], best1[:5])
        print('NCCL_SHARDED_OK reads=%%d ranks=%%d imbalance=%%.4f records=%%d hits=%%d' %% (N, world, loads.max() / loads.mean(), rec1.size + mrec1.size, n_hits1))
    else:
        assert all_rec is None and all_mrec is None
    eng.close()
    dist.barrier()
    dist.destroy_process_group()
''') % ROOT


def test_two_ranks_on_a_shard_reproduce_one_gpu(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two CUDA devices (gpurun --gpus 2)')
    script = tmp_path / 'worker.py'
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR='127.0.0.1')
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', '29541', str(script)],
                         capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    assert 'NCCL_SHARDED_OK' in out.stdout
    print(out.stdout.strip().splitlines()[-1])


PIPELINE_WORKER = textwrap.dedent('''
    import hashlib, os, sys, pathlib
    sys.path.insert(0, %r); sys.path.insert(0, os.path.join(%r, 'tests'))
    import torch
    import torch.distributed as dist
    import test_phases as tp
    from custom_porechop_abi_b200 import pipeline
    from custom_porechop_abi_b200.engine import AlignmentEngine
    local = int(os.environ['LOCAL_RANK'])
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    rank = dist.get_rank()
    tmp = pathlib.Path(sys.argv[1])
    eng = AlignmentEngine(local)
    for idx in (0, 5, 9, 12):
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
                                between_phases=between, engine=eng, scoring_scheme_vals=p['scoring'], end_size=p['end_size'],
                                adapter_threshold=p['adapter_threshold'], end_threshold=p['end_threshold'], extra_end_trim=p['extra_end_trim'],
                                min_trim_size=p['min_trim_size'], middle_threshold=p['middle_threshold'],
                                extra_middle_trim_good_side=p['good_side'], extra_middle_trim_bad_side=p['bad_side'],
                                discard_middle=a['discard_middle'], min_split_read_size=a['min_split'], barcode_threshold=p['barcode_threshold'],
                                barcode_diff=p['barcode_diff'], require_two_barcodes=c['require_two_barcodes'],
                                discard_unassigned=a['discard_unassigned'], untrimmed=a['untrimmed'], distributed=True, device='cuda')
            dist.barrier()
            if rank == 0:
                where = str(out / a['barcode_dir']) if a['barcode_dir'] else str(out)
                got = {fn: hashlib.sha256(open(os.path.join(where, fn), 'rb').read()).hexdigest() for fn in sorted(os.listdir(where))}
                assert got == g['files'][tag]['sha256'], (idx, tag, sorted(got), sorted(g['files'][tag]['sha256']))
    eng.close()
    if rank == 0:
        print('NCCL_PIPELINE_OK')
    dist.barrier()
    dist.destroy_process_group()
''') % (ROOT, ROOT)


def test_sharded_pipeline_on_two_gpus_writes_the_reference_bytes(tmp_path):
    """pipeline.trim on two ranks over NCCL, each with its shard of the reads on its own GPU: the files rank 0 writes are the
    ones the reference wrote (tests/golden/output_golden.json)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two CUDA devices (gpurun --gpus 2)')
    script = tmp_path / 'pipeline_worker.py'
    script.write_text(PIPELINE_WORKER)
    work = tmp_path / 'work'
    work.mkdir()
    env = dict(os.environ, MASTER_ADDR='127.0.0.1')
    out = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=2',
                          '--master-addr', '127.0.0.1', '--master-port', '29542', str(script), str(work)],
                         capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0, out.stderr[-3000:]
    assert 'NCCL_PIPELINE_OK' in out.stdout

