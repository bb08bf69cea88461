"""ctypes binding of libzebra_tppr.so (include/zebra_tppr.h).  No CPU fallback: if the library is
missing it is built; if it cannot be loaded, importing code gets the error."""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

_lib = None

_VP, _I64, _I32, _D = C.c_void_p, C.c_int64, C.c_int32, C.c_double

SIGNATURES = {
    "zebra_tppr_create": (C.c_int, [C.POINTER(_VP), _I64, _I32, _I32, _VP, _VP, _I32]),
    "zebra_tppr_destroy": (C.c_int, [_VP]),
    "zebra_tppr_table_bytes": (C.c_size_t, [_VP]),
    "zebra_tppr_record_bytes": (C.c_size_t, [_VP]),
    "zebra_tppr_bind_tables": (C.c_int, [_VP, _VP]),
    "zebra_tppr_reset_tickets": (C.c_int, [_VP, _VP]),
    "zebra_tppr_out_bytes": (C.c_size_t, [_VP, _I64, _I32]),
    "zebra_tppr_stream": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP, _VP, _VP]),
    "zebra_tppr_update_only": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _I64, _VP]),
    "zebra_tppr_plan_create": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I64, C.POINTER(_VP)]),
    "zebra_tppr_plan_destroy": (C.c_int, [_VP]),
    "zebra_tppr_plan_batches": (_I64, [_VP]),
    "zebra_tppr_plan_bytes": (C.c_size_t, [_VP]),
    "zebra_debug_host_pass": (C.c_int, [_I64, _I32, _I32, _VP, _VP, _VP, _VP, _VP, _I64, _I64, _VP, _VP, _VP, _VP, _VP]),
    "zebra_debug_read_record": (C.c_int, [_VP, _I32, _I64, _VP]),
    "zebra_debug_tie_order": (C.c_int, [_VP, _VP, _I32, _I32, _VP]),
    "zebra_tppr_plan_run": (C.c_int, [_VP, _VP, _I64, _I64, _I64, _VP, _I64, _VP]),
    "zebra_tppr_epoch_begin": (C.c_int, [_VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I64, _I64, _VP, C.POINTER(_VP)]),
    "zebra_tppr_epoch_next": (C.c_int, [_VP, C.POINTER(_VP), C.POINTER(_I64), C.POINTER(_I64)]),
    "zebra_tppr_epoch_end": (C.c_int, [_VP]),
    "zebra_tppr_epoch_slot_bytes": (_I64, [_VP]),
    "zebra_tppr_shard_init": (C.c_int, [_VP, _I32, _I32, _I64, C.c_size_t, _VP]),
    "zebra_tppr_shard_connect": (C.c_int, [_VP, _VP]),
    "zebra_tppr_shard_connect_local": (C.c_int, [_VP, _VP]),
    "zebra_tppr_shard_disconnect": (C.c_int, [_VP]),
    "zebra_tppr_shard_out": (_VP, [_VP]),
    "zebra_tppr_shard_out_bytes": (C.c_size_t, [_VP]),
    "zebra_tppr_shard_reset": (C.c_int, [_VP, _VP]),
    "zebra_tppr_shard_row_mode": (C.c_int, [_VP, _I32]),
    "zebra_tppr_set_tie_mode": (C.c_int, [_VP, _I32]),
    "zebra_tppr_set_max_blocks": (C.c_int, [_VP, _I32]),
    "zebra_tppr_read_state": (C.c_int, [_VP, _I32, _VP, _I64, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "zebra_tppr_sync_check": (C.c_int, [_VP, _VP]),
    "zebra_tppr_counters": (C.c_int, [_VP, _VP]),
    "zebra_tppr_profile": (C.c_int, [_VP, _VP]),
    "zebra_tppr_last_kernel_ms": (_D, [_VP]),
    "zebra_csr_create": (C.c_int, [C.POINTER(_VP), _VP, _VP, _VP, _VP, _I64, _I64, _I32]),
    "zebra_csr_destroy": (C.c_int, [_VP]),
    "zebra_prune_out_bytes": (C.c_size_t, [_I64, _I32]),
    "zebra_prune_topk": (C.c_int, [_VP, _VP, _VP, _I64, _I32, _I32, _D, _D, _I32, _VP, _VP, _VP]),
    "zebra_prune_launches": (_I64, [_VP]),
    "zebra_prune_topk_static": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, _VP, _I64, _I32, _I32, _D, _D, _I32, _VP, _VP]),
    "zebra_gather_fwd": (C.c_int, [_VP, _I32, _I64, _I32, _VP, _I64, _I32, _VP, _I64, _I32, _VP, _VP, _I32, _VP, _VP, _VP, _VP, _VP]),
    "zebra_gather_bwd": (C.c_int, [_VP, _I32, _I64, _I32, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP]),
    "zebra_gather_bwd_sorted": (C.c_int, [_I32, _I64, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP]),
    "zebra_gather_check": (C.c_int, [_VP]),
    "zebra_pending_select": (C.c_int, [_VP, _I32, _I64, _I32, _VP, _I64, _VP, _VP, _VP, _VP, _VP]),
    "zebra_gather_fwd_lazy": (C.c_int, [_VP, _I32, _I64, _I32, _VP, _I64, _I32, _VP, _I64, _I32, _VP, _VP, _I32, _VP, _VP, _VP, _VP,
                                        _VP, _VP, _VP, _VP]),
    "zebra_gather_bwd_lazy": (C.c_int, [_I32, _I64, _I32, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _I64, _I32, _I32, _VP]),
    "zebra_raw_messages": (C.c_int, [_VP, _VP, _VP, _VP, _I64, _VP, _I64, _I32, _VP, _VP, _I64, _I32, _VP, _VP, _I32, _VP, _VP]),
    "zebra_edge_metrics": (C.c_int, [_VP, _VP, _I64, _I64, _VP, _VP]),
    "zebra_last_error": (C.c_char_p, []),
    "zebra_build_info": (C.c_char_p, []),
}


def lib():
    global _lib
    if _lib is None:
        so = os.environ.get("ZEBRA_TPPR_LIB")                  # developer override (e.g. the -DZEBRA_PROFILE build)
        if not so:
            so = _build.build()                                # compiles only when csrc/ or include/ is newer than the .so
        L = C.CDLL(so)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(rc: int):
    if rc != 0:
        raise RuntimeError("libzebra_tppr: " + lib().zebra_last_error().decode())
