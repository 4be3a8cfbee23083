"""GPU: pcb_align_windows_static -- the entry point that allocates nothing (scratch in a file-scope device array) -- called
through the C ABI with device pointers, against pcb_align_windows on the same windows: every record bit-identical, on
calls in a row that leave each other's contents in the shared array (different adapters, kernel classes, window counts,
one call that needs several scratch batches).  Collected last on purpose."""
import ctypes
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DEFAULT = (3, -6, -5, -2)


def _windows(rng, adapter, n, stride, max_len):
    from oracle_lib import mutate
    out = []
    for _ in range(n):
        m = min(max_len, rng.choice([max_len] * 6 + [0, 1, 17, max_len - 1]))
        r = ''.join(rng.choice('ACGT') for _ in range(m))
        if m > 40 and rng.random() < 0.6:
            p = rng.randrange(-len(adapter) // 2, m - 10)
            ins = mutate(rng, adapter, rng.choice([0, 0.05, 0.15]))
            r = (ins[-p:] + r)[:m] if p < 0 else (r[:p] + ins + r[p:])[:m]
        if m > 20 and rng.random() < 0.1:
            q = rng.randrange(0, m - 5)
            r = r[:q] + 'NN' + r[q + 2:]
        out.append(r)
    return out


def _static_call(lib, torch, adapter, scores, d_buf, d_len, n, stride, d_out, stream=None):
    return lib.pcb_align_windows_static(0, adapter.encode(), len(adapter), scores[0], scores[1], scores[2], scores[3],
                                        ctypes.c_void_p(d_buf.data_ptr()), ctypes.c_void_p(d_len.data_ptr()), n, stride,
                                        ctypes.c_void_p(d_out.data_ptr()), ctypes.c_void_p(stream.cuda_stream if stream else 0))


def test_static_entry_point_matches_pcb_align_windows():
    import torch
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    from custom_porechop_abi_b200.engine import AlignmentEngine, RESULT_DTYPE
    lib = w.C_LIB
    rng = random.Random(2025)
    cases = []          # (adapter, scores, n_windows, stride, max window length)
    cases.append((''.join(rng.choice('ACGT') for _ in range(24)), DEFAULT, 700, 160, 150))
    cases.append((''.join(rng.choice('ACGT') for _ in range(102)), DEFAULT, 300, 160, 150))      # two-stripe class
    cases.append((''.join(rng.choice('ACGT') for _ in range(111)), (2, -3, -4, -1), 3000, 256, 250))   # several scratch batches
    cases.append((''.join(rng.choice('ACGT') for _ in range(31)), (1, -1, -3, -2), 8192, 160, 150))    # the most windows a call takes
    cases.append((''.join(rng.choice('ACGT') for _ in range(24)), DEFAULT, 1, 16, 9))
    eng = AlignmentEngine(0)
    side = torch.cuda.Stream()
    try:
        prepared = []
        for adapter, scores, n, stride, mlen in cases:
            wins = _windows(rng, adapter, n, stride, mlen)
            buf, lens = eng.pack_windows(wins, stride=stride)
            eng.set_scoring(scores)
            eng.set_adapters([adapter])
            want = eng.align_windows(buf, lens, [0]).reshape(-1).copy()
            d_buf = torch.from_numpy(np.ascontiguousarray(buf)).cuda()
            d_len = torch.from_numpy(np.ascontiguousarray(lens)).cuda()
            d_out = torch.full((n * RESULT_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device='cuda')
            prepared.append((adapter, scores, n, stride, d_buf, d_len, d_out, want))
        torch.cuda.synchronize()
        # calls in a row, twice over: the second round finds every other call's leftovers in the array
        for rnd in range(2):
            for k, (adapter, scores, n, stride, d_buf, d_len, d_out, want) in enumerate(prepared):
                d_out.fill_(0xAB)
                torch.cuda.synchronize()
                free0 = torch.cuda.mem_get_info()[0]
                st = side if (k + rnd) % 2 else None
                rc = _static_call(lib, torch, adapter, scores, d_buf, d_len, n, stride, d_out, st)
                assert rc == 0, lib.pcb_last_error(None)
                free1 = torch.cuda.mem_get_info()[0]
                if rnd == 1:            # kernels loaded, set-up done: the device's free memory does not move
                    assert free1 == free0, (k, free0, free1)
                got = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=RESULT_DTYPE)
                bad = [i for i in range(n) if got[i].tobytes() != want[i].tobytes()]
                assert not bad, (rnd, k, bad[:5], got[bad[0]], want[bad[0]])
        # spot check against the oracle as well (the existing entry point is itself tested against it elsewhere)
        from oracle_lib import oracle_align, rec_matches_oracle
        adapter, scores, n, stride, d_buf, d_len, d_out, want = prepared[1]
        host = d_buf.cpu().numpy().reshape(n, stride)
        lens = d_len.cpu().numpy()
        got = np.frombuffer(d_out.cpu().numpy().tobytes(), dtype=RESULT_DTYPE)
        for i in range(0, n, 7):
            read = host[i, :lens[i]].tobytes().decode()
            assert rec_matches_oracle(got[i], oracle_align(read, adapter, scores)), (read, adapter)
    finally:
        eng.close()


def test_static_entry_point_refuses_what_does_not_fit():
    import torch
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    lib = w.C_LIB
    d = torch.zeros(4096, dtype=torch.uint8, device='cuda')
    dl = torch.zeros(16, dtype=torch.int32, device='cuda')
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    call = lambda ad, sc, n, stride: lib.pcb_align_windows_static(0, ad.encode(), len(ad), sc[0], sc[1], sc[2], sc[3], p(d), p(dl), n, stride, p(d), None)
    PCB_ERR_RANGE = 3
    assert call('ACGT' * 6, DEFAULT, 8193, 160) == PCB_ERR_RANGE and b'8192' in lib.pcb_last_error(None)
    assert call('ACGT' * 6, DEFAULT, 4, 272) == PCB_ERR_RANGE and b'stride' in lib.pcb_last_error(None)
    assert call('ACGN' * 6, DEFAULT, 4, 160) == PCB_ERR_RANGE
    assert call('A' * 128, DEFAULT, 4, 160) == PCB_ERR_RANGE
    assert call('ACGT' * 6, (1, -1, -1, -1), 4, 160) == PCB_ERR_RANGE          # linear gaps: the general kernels' business
    assert call('ACGT' * 6, (3, -6, -9, -2), 4, 160) == PCB_ERR_RANGE          # gap open below -7
    # a window length outside [0, stride] is caught on the device and reported, nothing is written out of bounds
    dl[3] = 400
    rc = call('ACGT' * 6, DEFAULT, 16, 160)
    assert rc == 2 and b'window length' in lib.pcb_last_error(None)
    assert call('ACGT' * 6, DEFAULT, 0, 160) == 0
