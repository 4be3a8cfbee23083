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
