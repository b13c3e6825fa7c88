"""ctypes binding of ``csrc/libgemdat_b200.so`` (declared in ``include/gemdat_b200.h``).

The shared library is the product: there is no Python or CPU fallback.  Importing
this module without the built library raises, and every compute entry point
returns ``GD_ERR_CUDA`` when no CUDA device is usable.
"""

from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get('GEMDAT_B200_LIB', _HERE / 'csrc' / 'libgemdat_b200.so'))

GD_OK = 0
GD_ERR_CUDA = 1
GD_ERR_INVALID = 2
GD_ERR_ALIGNMENT = 3
GD_ERR_TRICLINIC = 4
GD_ERR_CAPACITY = 5
GD_ERR_RADIUS_GT_CUTOFF = 6
GD_ERR_INNER_CONSTANT = 7
GD_ERR_NO_EVENTS = 8
GD_ERR_NO_JUMPS = 9
GD_ERR_RANGE = 10
GD_JUMP_STATE_BYTES = 40
GD_JUMPS_TAIL_MISMATCH, GD_JUMPS_TAIL_UNPROVEN, GD_JUMPS_TAIL_UNRESOLVED, GD_JUMPS_TAIL_ROWS_OVERFLOW = 1, 2, 4, 8


class GemdatB200Error(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f'libgemdat_b200 error {code}: {message}')
        self.code = code
        self.message = message


def _require_library() -> None:
    if not LIB_PATH.exists():
        raise ImportError(
            f'{LIB_PATH} is missing: build it with `python -c "import __graft_entry__ as g; g.build()"` '
            'or gemdat_b200/csrc/build.sh (nvcc, sm_100a). gemdat_b200 has no CPU fallback.'
        )


_require_library()  # importing the package without the built library fails loudly


class _Library:
    """The shared library, mapped on first use: a process that only needs the pure-numpy helpers
    of the package (the synthetic recipe in bench.py's CPU reference arm) never loads it."""

    _cdll = None

    def load(self) -> C.CDLL:
        if self._cdll is None:
            _require_library()
            cdll = C.CDLL(str(LIB_PATH))
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(cdll, name)  # AttributeError here = the .so is stale: rebuild
                fn.restype = res
                fn.argtypes = args
            type(self)._cdll = cdll
        return self._cdll

    def __getattr__(self, name):
        return getattr(self.load(), name)


lib = _Library()

_p = C.c_void_p
_i32 = C.c_int32
_i64 = C.c_int64
_u64 = C.c_uint64
_sz = C.c_size_t

# name -> (restype, argtypes); must list every symbol include/gemdat_b200.h declares
SIGNATURES = {
    'gd_last_error': (C.c_char_p, []),
    'gd_version': (C.c_int, []),
    'gd_sm_count': (C.c_int, []),
    'gd_sitegrid_build': (C.c_int, [_p, _p, _p, _i32, _p, _p, _i32, _p, _sz, C.POINTER(_sz)]),
    'gd_site_states': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    'gd_events_stage_cap': (_i32, [_i64, _i32]),
    'gd_events_stage_elems': (_i64, [_i64, _i32, _i32]),
    'gd_site_states_staged': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i32, _p, _p]),
    'gd_events_compact': (C.c_int, [_p, _i32, _i64, _i32, _p, _p, _p, _i64, _p, _i64, _p]),
    'gd_remap_states': (C.c_int, [_p, _i64, _p, _i32, _p]),
    'gd_widen_i32_i64': (C.c_int, [_p, _p, _i64, _p]),
    'gd_events_work_elems': (_i64, [_i64, _i32]),
    'gd_events_count': (C.c_int, [_p, _p, _i64, _i32, _p, _p, _p, _p, _p]),
    'gd_events_scan': (C.c_int, [_p, _i64, _i32, _p, _p, C.POINTER(_i64), _p]),
    'gd_events_fill': (C.c_int, [_p, _p, _i64, _i32, _p, _p, _p, _p, _p, _i64, _p, _p]),
    'gd_events_merge': (C.c_int, [_p, _p, _p, _i32, _p, _p]),
    'gd_events_merge_capped': (C.c_int, [_p, _p, _p, _i32, _p, _i64, _i64, _p]),
    'gd_jumps_merge_shards': (C.c_int, [_p, _i64, _i32, _i32, _i32, _i64, _p, _i64, _p]),
    'gd_fill_states': (C.c_int, [_p, _i64, _i32, C.c_int, _p, _p, _p, _p, _p]),
    'gd_fill_states_noalloc': (C.c_int, [_p, _i64, _i32, C.c_int, _p, _p, _p, _p]),
    'gd_fill_last_valid': (C.c_int, [_p, _i64, _i32, C.c_int, _p, _p, _p]),
    'gd_fill_apply': (C.c_int, [_p, _i64, _i32, C.c_int, _p, _p, _p, _p]),
    'gd_occupancy': (C.c_int, [_p, _i64, _i32, _p, _p]),
    'gd_frame_mean': (C.c_int, [_p, _p, _p, _i64, _i32, _p, _p]),
    'gd_drift': (C.c_int, [_p, _i64, _i32, _p, _i32, _p, _p, _p]),
    'gd_cumsum_frames3_work_bytes': (_i64, [_i64]),
    'gd_cumsum_frames3': (C.c_int, [_p, _i64, _p, _p, _p]),
    'gd_drift_correct': (C.c_int, [_p, _p, _p, _i64, _i32, _p, _p]),
    'gd_residence_work_bytes': (_i64, [_i64]),
    'gd_residence_count': (C.c_int, [_p, _i64, _p, C.POINTER(_i64), _p]),
    'gd_residence_fill': (C.c_int, [_p, _i64, _p, _p, _p]),
    'gd_events_split_count': (C.c_int, [_p, _p, _i32, _p, _i32, _p, _p, _p]),
    'gd_events_split_fill': (C.c_int, [_p, _i32, _p, _i32, _p, _p, _p, _p, _p, _p]),
    'gd_jumps_work_bytes': (_i64, [_i64, _i32]),
    'gd_jumps_count': (C.c_int, [_p, _p, _i64, _i32, _i64, _i64, _p, C.POINTER(_i64), _p]),
    'gd_jumps_fill': (C.c_int, [_p, _p, _i64, _i32, _i64, _i64, _p, _p, _p]),
    'gd_jumps_fill_capped': (C.c_int, [_p, _p, _i64, _i32, _i64, _i64, _p, _p, _i64, _p]),
    'gd_jumps_count_shard': (C.c_int, [_p, _p, _i64, _i32, _i64, _i64, _p, C.c_int, C.c_int, _p, _p, _p, _p,
                                       C.POINTER(_i64), C.POINTER(_i32), _p]),
    'gd_jumps_work_unresolved_offset': (_i64, [_i64, _i32]),
    'gd_jumps_resolve': (C.c_int, [_p, _i32, _i32, _i32, _p, _p, C.POINTER(_i32), _p]),
    'gd_pair_matrix': (C.c_int, [_p, _i64, _i32, _i32, _i32, _i32, _p, _p, _p]),
    'gd_collective_work_bytes': (_i64, [_i64]),
    'gd_collective_sort': (C.c_int, [_p, _i64, _p, _p, _p]),
    'gd_collective_close': (C.c_int, [_p, _i32, C.c_double, _p, _p]),
    'gd_collective_count': (C.c_int, [_p, _i64, _p, _i32, _i32, _p, C.POINTER(_i64), C.POINTER(_i64), _p]),
    'gd_collective_fill': (C.c_int, [_p, _i64, _p, _i32, _i32, _p, _p, _p]),
    'gd_rdf_classes': (C.c_int, [_p, _p, _p, _i64, _i32, _p, _i32, _i32, _p, _p, _p, _p]),
    'gd_rdf_hist': (C.c_int, [_p, _i64, _i32, _p, _p, _i32, _p, _i32, _p, _p, _i32, _i32, _p, _i32,
                              C.c_double, _i32, _i32, _p, _p]),
    'gd_rdf_frames_work_bytes': (_i64, [_i64]),
    'gd_rdf_hist_frames': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _i32, _p, _p, _p, _i32, _i32, _p, _i32,
                                     C.c_double, _i32, _i32, _p, _p]),
    'gd_unwrap_chunks': (_i64, [_i64, _i32]),
    'gd_unwrap': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    'gd_unwrap_windings': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p]),
    'gd_unwrap_apply': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p]),
    'gd_amplitudes_work_bytes': (_i64, [_i64, _i32]),
    'gd_amplitudes_count': (C.c_int, [_p, _i64, _i32, _p, _p, C.POINTER(_i64), _p]),
    'gd_amplitudes_fill': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p]),
    'gd_mean_std_work_bytes': (_i64, []),
    'gd_mean_std': (C.c_int, [_p, _i64, _p, C.POINTER(C.c_double), C.POINTER(C.c_double), _p]),
    'gd_meanfreq': (C.c_int, [_p, _i64, _i32, C.c_double, _p, _p]),
    'gd_transpose_f64': (C.c_int, [_p, _i64, _i64, _p, _p]),
    'gd_wrap_unit': (C.c_int, [_p, _p, _i64, _p, _p]),
    'gd_transpose_soa': (C.c_int, [_p, _i64, _i32, _p, _p]),
    'gd_msd_direct': (C.c_int, [_p, _i64, _i32, _i64, _i64, _p, _p]),
    'gd_density': (C.c_int, [_p, _i64, _p, _p, _p, _p]),
    'gd_density_frames': (C.c_int, [_p, _i64, _i32, _p, _p, _p, _p]),
    'gd_lll_reduce_host': (C.c_int, [_p, _p, _p]),
    'gd_pbc_distances': (C.c_int, [_p, _p, _i32, _p, _i32, _p, _p]),
    'gd_synth_positions': (C.c_int, [_p, _p, _i32, _i32, _i32, _i32, _p, _u64, _i64, _i64, _p, _p]),
    'gd_fp64_peak': (C.c_int, [_i64, C.POINTER(C.c_double), _p]),
}

def last_error() -> str:
    return lib.gd_last_error().decode('utf-8', 'replace')


def check(status: int) -> None:
    """Raise for any non-zero status (callers map specific codes to the reference's exceptions first)."""
    if status != GD_OK:
        raise GemdatB200Error(status, last_error())
