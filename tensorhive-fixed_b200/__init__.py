"""B200-native reservation hot path of TensorHive (kivicode/TensorHive-Fixed v1.1.0).

Layers, bottom up:
  csrc/ + include/th_b200.h   hand-written sm_100a kernels behind a C ABI (libth_b200.so)
  native                      ctypes binding; fails loudly when the library or a GPU is missing
  api                         Snapshot / usage_reduce: numpy or torch-CUDA arrays in, arrays out
  calendar                    columnar snapshot builder (uuid dictionary, datetime -> int64 us, owner ids)
  dropin                      the reference's call surface (Reservation.would_interfere, current_events,
                              ProtectionService join, JobSchedulingService checks, avg) on top of it
  sharding                    uuid-range partition over ranks + torch.distributed all-gather of results
  synth                       synthetic calendars / usage logs of SURVEY.md section 8d

Import as ``tensorhive_fixed_b200`` (shim package at the repo root).
"""
from . import native  # noqa: F401
from .api import Snapshot, usage_reduce, usage_reduce_u8, sort_pairs, MODE_STRICT, MODE_INCLUSIVE, MODE_UPCOMING  # noqa: F401
from .api import OWNER_ANY, OWNER_OTHER, OWNER_SAME, NativeError  # noqa: F401

__all__ = ['native', 'Snapshot', 'usage_reduce', 'usage_reduce_u8', 'sort_pairs', 'MODE_STRICT', 'MODE_INCLUSIVE', 'MODE_UPCOMING',
           'OWNER_ANY', 'OWNER_OTHER', 'OWNER_SAME', 'NativeError']
