"""ctypes binding of libth_b200.so (declarations mirror include/th_b200.h one to one).

There is no CPU fallback: if the library is missing, or a compute call is made on a machine without a
B200, the call raises.  ``load()`` alone needs no GPU (the CPU test-suite uses it to check the exports).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# TH_B200_LIB (diagnostics only): load another build of the same library, e.g. an experiment variant under lib/variants/
LIB_PATH = os.environ.get('TH_B200_LIB') or os.path.join(_HERE, 'lib', 'libth_b200.so')

MEM_HOST, MEM_DEVICE = 0, 1

_i64p = C.POINTER(C.c_int64)
_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_vp = C.c_void_p
_i64, _int = C.c_int64, C.c_int

# name -> (restype, argtypes); every symbol include/th_b200.h declares
SIGNATURES = {
    'th_version': (_int, []),
    'th_last_error': (C.c_char_p, []),
    'th_pool_stats': (_int, [_int, _i64p, _i64p, _i64p, _int]),
    'th_snapshot_create': (_int, [_int, C.POINTER(_vp)]),
    'th_snapshot_destroy': (_int, [_vp]),
    'th_snapshot_load': (_int, [_vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _int, _vp]),
    'th_snapshot_apply_delta': (_int, [_vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _int, _vp]),
    'th_snapshot_info': (_int, [_vp, _i64p, _i64p, _i64p, _i64p]),
    'th_snapshot_export': (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _int, _vp]),
    'th_conflict_bitmap': (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _int, _vp, _int, _vp]),
    'th_conflict_bitmap_packed': (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _int, _vp, _int, _vp]),
    'th_conflict_sets': (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _int, _vp, _vp, _vp, _i64, _i64p, _int, _vp]),
    'th_current_owner': (_int, [_vp, _i64, _vp, _vp, _int, _vp]),
    'th_upcoming': (_int, [_vp, _i64, _i64, _i64, _vp, _vp, _vp, _vp, _vp, _int, _vp]),
    'th_upcoming_owner_flags': (_int, [_vp, _i64, _i64, _i64, _vp, _vp, _int, _vp, _int, _vp]),
    'th_violation_join': (_int, [_vp, _i64, _int, _i64, _vp, _vp, _vp, _vp, _int, _vp]),
    'th_violation_join_static': (_int, [_vp, _i64, _int, _i64, _vp, _vp, _vp, _vp, _vp]),
    'th_usage_reduce': (_int, [_int, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                               _int, _vp]),
    'th_peer_create': (_int, [_int, _int, _int, _i64, C.POINTER(_vp), _vp]),
    'th_peer_connect': (_int, [_vp, _vp, _vp]),
    'th_peer_local_base': (_int, [_vp, C.POINTER(_vp)]),
    'th_peer_destroy': (_int, [_vp]),
    'th_peer_check': (_int, [_vp, _vp]),
    'th_conflict_bitmap_exchange': (_int, [_vp, _vp, _i64, _vp, _vp, _vp, _vp, _int, _vp, _int, C.POINTER(_vp), _vp]),
    'th_conflict_bitmap_exchange_packed': (_int, [_vp, _vp, _i64, _vp, _vp, _vp, _vp, _int, _vp, _int, C.POINTER(_vp), _vp]),
    'th_peer_wait': (_int, [_vp, _vp]),
    'th_peer_barrier': (_int, [_vp, _vp]),
    'th_usage_reduce_u8': (_int, [_int, _i64, _i64, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp,
                                  _int, _vp]),
    'th_ring_create': (_int, [_int, _i64, C.POINTER(_vp)]),
    'th_ring_destroy': (_int, [_vp]),
    'th_ring_append': (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp]),
    'th_ring_reduce': (_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _int, _vp]),
    'th_ring_info': (_int, [_vp, _i64p, _i64p, _i64p, _i64p, _i64p, _i64p]),
    'th_sort_pairs': (_int, [_int, _i64, _vp, _vp, _vp, _int, _vp]),
    'th_synth_samples': (_int, [_int, _i64, _i64, _i64, C.c_uint64, _vp, _vp, _vp, _vp, _vp]),
    'th_synth_windows': (_int, [_int, _i64, _i64, _i64, C.c_uint64, _i64, _vp, _vp, _vp, _vp]),
}

_lib = None


class NativeError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__('libth_b200 error %d: %s' % (code, message))
        self.code = code


def load() -> C.CDLL:
    """dlopen the in-tree library and bind every declared symbol.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError('%s is missing: run `python __graft_entry__.py` (build()) first; there is no CPU '
                          'fallback for this path' % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the export is missing
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error() -> str:
    return load().th_last_error().decode('utf-8', 'replace')


def check(rc: int) -> int:
    if rc != 0:
        raise NativeError(rc, last_error())
    return rc
