"""
Multi-GPU plumbing for the hot path: reads are independent, so ranks take length-balanced shards and never
exchange data during the DP.  Two tiny merges exist, both mirroring reductions the reference does in-process:

* phase A keeps, per adapter set, the MAX full-adapter identity over all checked reads
  (nanopore_read.py:168,173 -> adapters.py best_start_score / best_end_score, thresholded at
  porechop_abi.py:245): ranks all-reduce(MAX) a float64 vector of 2 x n_sets entries;
* the summaries count reads with start / end / middle hits (porechop_abi.py:441-454, :525-532):
  all-reduce(SUM) of a handful of int64 counters.

Backend agnostic (NCCL on the GPU box, gloo in the CPU tests); with no process group it is the identity.
"""
import numpy as np


def balanced_shards(lengths, world_size):
    """
    Longest-processing-time partition of reads by length: returns a list (one entry per rank) of sorted index
    arrays.  Work per read is proportional to its length (middle search) plus a constant (end windows), so the
    weight is length + window_cost.
    """
    lengths = np.asarray(lengths, dtype=np.int64)
    order = np.argsort(-lengths, kind='stable')
    loads = np.zeros(world_size, dtype=np.int64)
    owner = np.empty(len(lengths), dtype=np.int32)
    window_cost = 2 * 150 * 59          # 2 windows x 236 adapter sequences / 4 middle adapters, in read bases
    # greedy LPT; vectorised in blocks of world_size (reads sorted by length are near-equal inside a block)
    for start in range(0, len(order), world_size):
        block = order[start:start + world_size]
        ranks = np.argsort(loads, kind='stable')[:len(block)]
        owner[block] = ranks
        loads[ranks] += lengths[block] + window_cost
    return [np.sort(np.nonzero(owner == r)[0]) for r in range(world_size)]


def bind_to_gpu_cpus(device_index):
    """One process per GPU: run this process (and allocate its pinned buffers, first touch) on the CPUs next to its GPU.
    With 8 ranks on one host the copies of the host-buffer calls otherwise cross the socket interconnect.  Returns the CPU
    list it bound to, or None when NVML or the affinity call is not available (nothing is changed then)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        handle = pynvml.nvmlDeviceGetHandleByIndex(int(device_index))
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        cpus = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1 and 64 * w + b < n_cpu]
        allowed = os.sched_getaffinity(0)
        cpus = [c for c in cpus if c in allowed]
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:                                       # no NVML / no permission: stay where the launcher put us
        return None


def _dist():
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist
    except Exception:
        pass
    return None


def merge_best_identities(best, device=None):
    """all-reduce(MAX) of the per-(set, end) best full-adapter identities (NaN never wins, as Python's max)."""
    import torch
    dist = _dist()
    best = np.asarray(best, dtype=np.float64)
    if dist is None:
        return best
    t = torch.from_numpy(np.nan_to_num(best, nan=-1.0)).to(device or 'cpu')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.cpu().numpy()


def merge_counters(counts, device=None):
    """all-reduce(SUM) of int64 summary counters."""
    import torch
    dist = _dist()
    counts = np.asarray(counts, dtype=np.int64)
    if dist is None:
        return counts
    t = torch.from_numpy(counts.copy()).to(device or 'cpu')
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.cpu().numpy()


def gather_results(local_indices, local_records, n_total, device=None):
    """
    Rank 0 receives every rank's records in global read order (the only "collective" the data path has:
    gathering per-read results).  local_records: numpy array whose first axis matches local_indices.
    Returns the merged array on rank 0 and None elsewhere; identity without a process group.
    """
    dist = _dist()
    if dist is None:
        out = np.empty((n_total,) + local_records.shape[1:], dtype=local_records.dtype)
        out[local_indices] = local_records
        return out
    gathered = [None] * dist.get_world_size()
    dist.all_gather_object(gathered, (np.asarray(local_indices), local_records))
    if dist.get_rank() != 0:
        return None
    out = np.empty((n_total,) + local_records.shape[1:], dtype=local_records.dtype)
    for idx, rec in gathered:
        out[idx] = rec
    return out
