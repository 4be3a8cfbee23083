"""
Derived from Porechop / Porechop_ABI's porechop_abi/cpp_function_wrappers.py (Copyright 2017 Ryan Wick; Porechop_ABI by
Quentin Bonenfant et al.), which is distributed under the GNU General Public License v3 or later; the argtypes block of
adapterAlignment and the two functions adapter_alignment / c_string_to_python_string below follow that file, because they
ARE the drop-in boundary.  This file is distributed under the same terms.  See <http://www.gnu.org/licenses/>.

ctypes binding of the B200 adapter-alignment library -- the drop-in for the reference's
``porechop_abi/cpp_function_wrappers.py`` (reference lines 21-63).

Same module-level contract as the reference:

* the shared library is ``cpp_functions.so`` next to this file and is loaded at import time; a missing
  library is a hard error (reference :21-25) -- there is no CPU fallback of any kind;
* ``adapter_alignment(read_sequence, adapter_sequence, scoring_scheme_vals) -> str`` with
  ``scoring_scheme_vals = [match, mismatch, gap_open, gap_extend]`` (reference :42-53) returns the
  seven-field comma separated string of ``ScoredAlignment::getString`` (src/alignment.cpp:112-120);
* ``c_string_to_python_string`` frees the C string with ``freeCString`` (reference :56-63).

On top of that, ``AlignmentEngine`` exposes the batched entry points of include/porechop_b200.h: the
reference calls ``adapter_alignment`` once per (window, adapter) pair from a thread pool
(porechop_abi.py:228-239, :406-432, :500-517); a GPU wants all pairs of a phase in one submit.
"""
import ctypes
import os
import sys
from ctypes import CDLL, POINTER, Structure, byref, c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p, cast

SO_FILE = 'cpp_functions.so'
SO_FILE_FULL = os.environ.get('PCB_SO') or os.path.join(os.path.dirname(os.path.realpath(__file__)), SO_FILE)
if not os.path.isfile(SO_FILE_FULL):
    sys.exit('could not find ' + SO_FILE + ' - build it with: python -c "import __graft_entry__ as g; g.build()"')
C_LIB = CDLL(SO_FILE_FULL)

C_LIB.adapterAlignment.argtypes = [c_char_p,  # Read sequence
                                   c_char_p,  # Adapter sequence
                                   c_int,     # Match score
                                   c_int,     # Mismatch score
                                   c_int,     # Gap open score
                                   c_int]     # Gap extension score
C_LIB.adapterAlignment.restype = c_void_p     # String describing alignment
C_LIB.freeCString.argtypes = [c_void_p]
C_LIB.freeCString.restype = None
C_LIB.pcb_set_dropin_fatal.argtypes = [c_int]
C_LIB.pcb_set_dropin_fatal(0)                 # this wrapper checks pcb_last_error itself and raises (adapter_alignment below)


class AlignmentError(RuntimeError):
    """Raised when the CUDA library reports an error (no device, out of range scoring scheme ...)."""


class PcbResult(Structure):
    _fields_ = [('read_start', c_int32), ('read_end', c_int32), ('adapter_start', c_int32),
                ('adapter_end', c_int32), ('raw_score', c_int32), ('matches', c_int32),
                ('aligned_len', c_int32), ('full_len', c_int32)]


class PcbBest(Structure):
    _fields_ = [('matches', c_int32), ('full_len', c_int32)]


class PcbHitFilter(Structure):
    _fields_ = [('threshold_e6', c_int64), ('at_end', c_int32), ('end_size', c_int32), ('min_trim_size', c_int32),
                ('reserved', c_int32)]


class PcbHit(Structure):
    _fields_ = [('window', c_int32), ('id_pos', c_int32), ('read_start', c_int32), ('read_end', c_int32),
                ('raw_score', c_int32), ('matches', c_int32), ('aligned_len', c_int32), ('full_len', c_int32)]


class PcbCounters(Structure):
    _fields_ = [('kernel_launches', c_int64), ('dp_cells', c_int64), ('h2d_bytes', c_int64),
                ('d2h_bytes', c_int64), ('dp_kernel_launches', c_int64), ('dp_kernel_ms', c_float)]


_P = c_void_p
C_LIB.pcb_create.argtypes = [c_int, POINTER(_P)]
C_LIB.pcb_destroy.argtypes = [_P]
C_LIB.pcb_destroy.restype = None
C_LIB.pcb_last_error.argtypes = [_P]
C_LIB.pcb_last_error.restype = c_char_p
C_LIB.pcb_device_count.argtypes = []
C_LIB.pcb_select_slot.argtypes = [_P, c_int]
C_LIB.pcb_set_scoring.argtypes = [_P, c_int, c_int, c_int, c_int]
C_LIB.pcb_set_adapters.argtypes = [_P, c_char_p, POINTER(c_int64), c_int32]
C_LIB.pcb_align_windows.argtypes = [_P, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int32, c_void_p]
C_LIB.pcb_align_windows_best.argtypes = [_P, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int32, c_void_p]
C_LIB.pcb_align_windows_hits.argtypes = [_P, c_void_p, c_int64, c_int32, c_void_p, c_void_p, c_int32, POINTER(PcbHitFilter),
                                         c_void_p, c_void_p, c_int64, POINTER(c_int64), c_void_p]
C_LIB.check_compatibility.argtypes = [c_char_p, c_char_p]
C_LIB.check_compatibility.restype = c_int
C_LIB.pcb_check_compatibility.argtypes = [_P, c_char_p, POINTER(c_int64), c_int32, c_void_p, c_void_p, c_int64, c_void_p]
C_LIB.pcb_fetch_top2.argtypes = [_P, c_void_p, c_int64]
C_LIB.pcb_fetch_trims.argtypes = [_P, c_void_p, c_int64]
C_LIB.pcb_align_tasks.argtypes = [_P, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64, c_void_p]
C_LIB.pcb_upload_reads.argtypes = [_P, c_void_p, c_int64]
C_LIB.pcb_set_deferred.argtypes = [_P, c_int]
C_LIB.pcb_wait.argtypes = [_P]
C_LIB.pcb_wait_slot.argtypes = [_P, c_int]
C_LIB.pcb_stage_windows.argtypes = [_P, c_void_p, c_int64, c_int32, c_void_p]
C_LIB.pcb_stage_windows_from_reads.argtypes = [_P, c_void_p, c_void_p, c_int64, c_int32, c_int32]
C_LIB.pcb_stage_ids.argtypes = [_P, c_void_p, c_int32]
C_LIB.pcb_run_windows_resident.argtypes = [_P]
C_LIB.pcb_fetch_results.argtypes = [_P, c_void_p, c_int64]
C_LIB.pcb_stage_tasks.argtypes = [_P, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int64]
C_LIB.pcb_run_tasks_resident.argtypes = [_P]
C_LIB.pcb_timer_start.argtypes = [_P]
C_LIB.pcb_timer_stop.argtypes = [_P, POINTER(c_float)]
C_LIB.pcb_synchronize.argtypes = [_P]
C_LIB.pcb_get_counters.argtypes = [_P, POINTER(PcbCounters)]
C_LIB.pcb_reset_counters.argtypes = [_P]
C_LIB.pcb_set_kernel_timing.argtypes = [_P, c_int]
C_LIB.pcb_format_result.argtypes = [POINTER(PcbResult), c_char_p, c_uint64]
C_LIB.pcb_format_result.restype = None
C_LIB.pcb_stats_from_path.argtypes = [c_char_p, c_int64, c_int64, c_int64, c_int64, c_int64, c_char_p, c_int32,
                                      POINTER(PcbResult)]
# the entry point that allocates nothing (device pointers in and out; include/porechop_b200.h)
C_LIB.pcb_align_windows_static.argtypes = [c_int, c_char_p, c_int32, c_int, c_int, c_int, c_int, c_void_p, c_void_p,
                                           c_int64, c_int32, c_void_p, c_void_p]
C_LIB.pcb_align_windows_static.restype = c_int


def adapter_alignment(read_sequence, adapter_sequence, scoring_scheme_vals):
    """
    Python wrapper for the adapterAlignment C function (reference cpp_function_wrappers.py:42-53).
    Runs on the GPU; raises AlignmentError if the library could not run it.
    """
    match_score = scoring_scheme_vals[0]
    mismatch_score = scoring_scheme_vals[1]
    gap_open_score = scoring_scheme_vals[2]
    gap_extend_score = scoring_scheme_vals[3]
    ptr = C_LIB.adapterAlignment(read_sequence.encode('utf-8'), adapter_sequence.encode('utf-8'),
                                 match_score, mismatch_score, gap_open_score, gap_extend_score)
    result_string = c_string_to_python_string(ptr)
    err = C_LIB.pcb_last_error(None)
    if err:
        raise AlignmentError(err.decode())
    return result_string


def c_string_to_python_string(c_string):
    """
    Casts a C string to a Python string and frees the C string (reference cpp_function_wrappers.py:56-63).
    """
    python_string = cast(c_string, c_char_p).value.decode()
    C_LIB.freeCString(c_string)
    return python_string


def check_compatibility(seq1, seq2):
    """compatibility.so's check_compatibility (consensus.py:72-100 binds it the same way): 0 / 1 / 2."""
    flag = C_LIB.check_compatibility(seq1.encode('utf-8'), seq2.encode('utf-8'))
    err = C_LIB.pcb_last_error(None)
    if err:
        raise AlignmentError(err.decode())
    return flag


def format_result(rec):
    """The reference's result string for one PcbResult record (or numpy record with the same fields)."""
    r = rec if isinstance(rec, PcbResult) else PcbResult(*[int(rec[f]) for f, _ in PcbResult._fields_])
    buf = ctypes.create_string_buffer(192)
    C_LIB.pcb_format_result(byref(r), buf, 192)
    return buf.value.decode()
