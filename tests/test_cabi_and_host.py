"""CPU: the C-ABI library loads, exports every symbol include/porechop_b200.h declares, refuses to compute without
a device (no CPU fallback), and the host-side helpers behave."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, 'include', 'porechop_b200.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(pcb_[a-z_]+|adapterAlignment|freeCString)\s*\(', text)))


def test_library_exports_every_declared_symbol():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    syms = declared_symbols()
    assert 'adapterAlignment' in syms and 'freeCString' in syms and len(syms) >= 25
    for s in syms:
        assert hasattr(w.C_LIB, s), s


def test_static_entry_point_checks_its_limits_before_it_touches_a_device():
    """pcb_align_windows_static refuses what its fixed scratch array cannot hold with PCB_ERR_RANGE and a text, on the host,
    before any CUDA call (so this runs without a GPU; the compute side is tests/test_zz_no_alloc.py)."""
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    lib = w.C_LIB
    fake = ctypes.c_void_p(4096)                    # never dereferenced on these paths
    call = lambda ad, sc, n, stride: lib.pcb_align_windows_static(0, ad.encode(), len(ad), sc[0], sc[1], sc[2], sc[3], fake, fake, n, stride, fake, None)
    d = (3, -6, -5, -2)
    assert call('ACGT' * 6, d, 8193, 160) == 3 and b'8192' in lib.pcb_last_error(None)
    assert call('ACGT' * 6, d, 16, 257) == 3 and b'stride' in lib.pcb_last_error(None)
    assert call('A' * 128, d, 16, 160) == 3 and b'127' in lib.pcb_last_error(None)
    assert call('ACGN', d, 16, 160) == 3
    assert call('ACGT' * 6, (2, -3, -4, -4), 16, 160) == 3 and b'affine' in lib.pcb_last_error(None)
    assert call('ACGT' * 6, d, 0, 160) == 0        # nothing to do is not an error
    assert lib.pcb_align_windows_static(0, None, 0, 3, -6, -5, -2, fake, fake, 4, 160, fake, None) == 2


def test_result_record_layout():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    from custom_porechop_abi_b200.engine import RESULT_DTYPE
    assert ctypes.sizeof(w.PcbResult) == 32 == RESULT_DTYPE.itemsize
    assert [f for f, _ in w.PcbResult._fields_] == list(RESULT_DTYPE.names)


def test_format_result_matches_reference_print():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    r = w.PcbResult(8, 14, 0, 6, 21, 7, 7, 12)
    assert w.format_result(r) == '8,14,0,6,21,100.000000,58.333333'
    r = w.PcbResult(0, 0, 7, 6, 0, 0, 0, 7)
    assert w.format_result(r) == '0,0,7,6,0,-nan,0.000000'      # 0.0 / 0 as the reference prints it
    r = w.PcbResult(-1, 0, -1, 0, -2147483648, 0, 0, 0)
    assert w.format_result(r) == '-1,0,-1,0,-2147483648,0.000000,0.000000'


def _no_gpu():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    return w.C_LIB.pcb_device_count() == 0


@pytest.mark.skipif(not _no_gpu(), reason='a CUDA device is present')
def test_no_cpu_fallback_without_device():
    from custom_porechop_abi_b200 import cpp_function_wrappers as w
    from custom_porechop_abi_b200.engine import AlignmentEngine
    with pytest.raises(w.AlignmentError):
        AlignmentEngine(0)
    with pytest.raises(w.AlignmentError):
        w.adapter_alignment('ACGTACGT', 'ACGT', [3, -6, -5, -2])


def test_balanced_shards_cover_and_balance():
    from custom_porechop_abi_b200.sharding import balanced_shards
    rng = np.random.default_rng(0)
    lengths = np.clip(rng.lognormal(8.75, 0.75, 5000), 200, 200000).astype(np.int64)
    for world in (1, 2, 4, 8):
        shards = balanced_shards(lengths, world)
        allidx = np.sort(np.concatenate(shards))
        assert np.array_equal(allidx, np.arange(len(lengths)))
        loads = np.array([lengths[s].sum() for s in shards], dtype=np.float64)
        assert loads.max() / loads.mean() < 1.02


def test_synthetic_workload_shape():
    from custom_porechop_abi_b200 import synthetic
    adapters, s_ids, e_ids = synthetic.adapter_set()
    assert len(s_ids) == 119 and len(e_ids) == 117 and sum(len(a) for a in adapters) == 5747
    blob, off = synthetic.make_reads(200, adapters, s_ids, e_ids, seed=5)
    sbuf, ebuf, lens = synthetic.end_windows(blob, off, 150, 160)
    for r in (0, 17, 199):
        read = bytes(blob[off[r]:off[r + 1]])
        assert bytes(sbuf[r, :lens[r]]) == read[:150] and bytes(ebuf[r, :lens[r]]) == read[-150:]


def test_no_oracle_in_product():
    """The product path may not reach the oracle (or any CPU alignment code): nothing under the package, the public
    header or the C sources names oracle/ or its symbols."""
    bad = []
    for base in (os.path.join(ROOT, 'custom_porechop_abi_b200'), os.path.join(ROOT, 'include')):
        for dp, _, files in os.walk(base):
            for f in files:
                if f.endswith(('.py', '.cu', '.h', '.cpp', '.c')):
                    text = open(os.path.join(dp, f), errors='replace').read()
                    if re.search(r'oracle|pc_oracle_|libporechop_oracle|cpp_functions_ref', text):
                        bad.append(os.path.join(dp, f))
    assert not bad, bad
