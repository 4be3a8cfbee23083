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
    """all-reduce(MAX) of the per-(set, end) best full-adapter identities.  A NaN never wins, as in the reference's
    max(best, score) (nanopore_read.py:168,173), where best starts at 0.0."""
    import torch
    dist = _dist()
    best = np.nan_to_num(np.asarray(best, dtype=np.float64), nan=0.0)
    if dist is None:
        return best
    t = torch.from_numpy(best.copy()).to(device or 'cpu')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.cpu().numpy()


def merge_best(best, device=None):
    """Phase A across ranks, exactly: `best` is what pcb_align_windows_best returns on this rank (matches, full_len per
    adapter sequence; (0, 0) = nothing aligned).  Every rank gets, per sequence, the pair with the largest
    matches / full_len over all ranks (compared by cross-multiplication, so no float is involved before the caller
    prints it the way the reference's library does)."""
    import torch
    dist = _dist()
    pairs = np.stack([np.asarray(best['matches'], dtype=np.int64), np.asarray(best['full_len'], dtype=np.int64)], axis=1)
    if dist is None:
        return pairs
    t = torch.from_numpy(pairs.copy()).to(device or 'cpu')
    parts = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(parts, t)
    out = parts[0].cpu().numpy().copy()
    for q in parts[1:]:
        q = q.cpu().numpy()
        better = q[:, 0] * np.maximum(out[:, 1], 1) * (q[:, 1] > 0) > out[:, 0] * np.maximum(q[:, 1], 1) * (out[:, 1] > 0)
        out[better] = q[better]
    return out


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
    gathering per-read results).  local_records: numpy array (structured dtypes welcome) whose first axis matches
    local_indices.  Returns the merged array on rank 0 and None elsewhere; identity without a process group.
    The records travel as raw bytes in tensors (a gather to rank 0, padded to the largest shard), not as pickles.
    """
    import torch
    dist = _dist()
    local_indices = np.asarray(local_indices, dtype=np.int64)
    local_records = np.ascontiguousarray(local_records)
    if dist is None:
        out = np.empty((n_total,) + local_records.shape[1:], dtype=local_records.dtype)
        out[local_indices] = local_records
        return out
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = device or 'cpu'
    row_bytes = local_records.dtype.itemsize * int(np.prod(local_records.shape[1:], dtype=np.int64))
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    counts[rank] = len(local_indices)
    dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    counts = counts.cpu().numpy()
    cap = int(counts.max())
    idx = torch.zeros(cap, dtype=torch.int64, device=dev)
    idx[:len(local_indices)] = torch.from_numpy(local_indices).to(dev)
    payload = torch.zeros(cap * row_bytes, dtype=torch.uint8, device=dev)
    raw = np.frombuffer(local_records.tobytes(), dtype=np.uint8)
    payload[:len(raw)] = torch.from_numpy(raw.copy()).to(dev)
    idx_parts = [torch.empty_like(idx) for _ in range(world)] if rank == 0 else None
    pay_parts = [torch.empty_like(payload) for _ in range(world)] if rank == 0 else None
    dist.gather(idx, idx_parts, dst=0)
    dist.gather(payload, pay_parts, dst=0)
    if rank != 0:
        return None
    out = np.empty((n_total,) + local_records.shape[1:], dtype=local_records.dtype)
    for r in range(world):
        k = int(counts[r])
        rec = np.frombuffer(pay_parts[r][:k * row_bytes].cpu().numpy().tobytes(), dtype=local_records.dtype)
        out[idx_parts[r][:k].cpu().numpy()] = rec.reshape((k,) + local_records.shape[1:])
    return out


def gather_records(local_indices, local_records, device=None):
    """
    The output side of a sharded run: every rank hands in what read_output.records_of made of ITS reads -- per kept read
    (global read index, (bin, text, bases)) -- and rank 0 gets all of them in global read order (what one process would have
    written); None elsewhere; identity without a process group.  Bytes and int64 arrays in tensors, no pickles.
    local_indices[i] belongs to local_records[i]; a read with several records appears several times.
    """
    import torch
    dist = _dist()
    order = np.argsort(np.asarray(local_indices, dtype=np.int64), kind='stable')
    if dist is None:
        return [local_records[i] for i in order]
    world, rank = dist.get_world_size(), dist.get_rank()
    dev = device or 'cpu'
    names = [local_records[i][0].encode() for i in range(len(local_records))]
    texts = [local_records[i][1].encode() for i in range(len(local_records))]
    meta = np.zeros((len(local_records), 4), dtype=np.int64)           # read index, bases, bytes of the bin's name, bytes of the text
    if len(local_records):
        meta[:, 0] = np.asarray(local_indices, dtype=np.int64)
        meta[:, 1] = [r[2] for r in local_records]
        meta[:, 2] = [len(x) for x in names]
        meta[:, 3] = [len(x) for x in texts]
    payload = np.frombuffer(b''.join(a + b for a, b in zip(names, texts)), dtype=np.uint8)
    sizes = torch.zeros(world, 2, dtype=torch.int64, device=dev)
    sizes[rank, 0], sizes[rank, 1] = len(meta), len(payload)
    dist.all_reduce(sizes, op=dist.ReduceOp.SUM)
    sizes = sizes.cpu().numpy()
    m_max, p_max = int(sizes[:, 0].max()), int(sizes[:, 1].max())
    m_send = torch.zeros(max(m_max, 1), 4, dtype=torch.int64, device=dev)
    p_send = torch.zeros(max(p_max, 1), dtype=torch.uint8, device=dev)
    if len(meta):
        m_send[:len(meta)] = torch.from_numpy(meta).to(dev)
    if len(payload):
        p_send[:len(payload)] = torch.from_numpy(payload.copy()).to(dev)
    m_all = [torch.empty_like(m_send) for _ in range(world)]
    p_all = [torch.empty_like(p_send) for _ in range(world)]
    dist.all_gather(m_all, m_send)
    dist.all_gather(p_all, p_send)
    if rank != 0:
        return None
    merged = []
    for r in range(world):
        m = m_all[r].cpu().numpy()[:int(sizes[r, 0])]
        p = p_all[r].cpu().numpy()[:int(sizes[r, 1])].tobytes()
        at = 0
        for idx, bases, ln, lt in m.tolist():
            merged.append((idx, len(merged), p[at:at + ln].decode(), p[at + ln:at + ln + lt].decode(), bases))
            at += ln + lt
    merged.sort(key=lambda x: (x[0], x[1]))                              # read order; a read's own records keep theirs
    return [(name, text, bases) for _, _, name, text, bases in merged]
