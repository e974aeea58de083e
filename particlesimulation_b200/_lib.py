"""ctypes binding of libparsim_b200.so (the C ABI in include/parsim_b200.h).

There is no CPU fallback: if the shared library has not been built the import fails loudly,
and every compute entry point fails with the library's own error when no CUDA device exists.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PARSIM_LIB") or os.path.join(HERE, "libparsim_b200.so")  # PARSIM_LIB: tuning builds
NKERNELS = 16


class Stats(C.Structure):
    _fields_ = [
        ("steps", C.c_longlong), ("particle_slots", C.c_longlong),
        ("alive_particle_steps", C.c_longlong), ("force_pairs", C.c_longlong),
        ("collision_pairs", C.c_longlong), ("movers", C.c_longlong), ("strays", C.c_longlong),
        ("deaths", C.c_longlong), ("migrants", C.c_longlong), ("run_ms", C.c_double),
        ("kernel_ms", C.c_double * NKERNELS), ("kernel_launches", C.c_longlong * NKERNELS),
        ("kernel_bytes", C.c_double * NKERNELS),
    ]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_ctx = C.c_void_p

# name -> (restype, argtypes); must list every symbol include/parsim_b200.h declares
SIGNATURES = {
    "parsim_abi_version": (C.c_int, []),
    "parsim_create": (C.c_int, [C.POINTER(_ctx), C.c_double, C.c_long, C.c_longlong, C.c_int]),
    "parsim_nccl_unique_id": (C.c_int, [C.POINTER(C.c_ubyte)]),
    "parsim_create_rank": (C.c_int, [C.POINTER(_ctx), C.c_double, C.c_long, C.c_longlong, C.c_int,
                                     C.c_int, C.c_int, C.POINTER(C.c_ubyte)]),
    "parsim_destroy": (None, [_ctx]),
    "parsim_last_error": (C.c_char_p, [_ctx]),
    "parsim_upload": (C.c_int, [_ctx, _dp, _dp, _dp, _dp, _dp]),
    "parsim_upload_begin": (C.c_int, [_ctx]),
    "parsim_upload_chunk": (C.c_int, [_ctx, C.c_longlong, C.c_longlong, _dp, _dp, _dp, _dp, _dp]),
    "parsim_upload_end": (C.c_int, [_ctx]),
    "parsim_upload_chunk_ids": (C.c_int, [_ctx, C.c_longlong, C.POINTER(C.c_uint), _dp, _dp, _dp, _dp, _dp]),
    "parsim_upload_strip": (C.c_int, [_ctx, C.c_longlong, C.POINTER(C.c_uint), _dp, _dp, _dp, _dp, _dp]),
    "parsim_download": (C.c_int, [_ctx, C.c_longlong, C.POINTER(C.c_uint), _dp, _dp, _dp, _dp, _dp,
                                  C.POINTER(C.c_longlong)]),
    "parsim_reset": (C.c_int, [_ctx]),
    "parsim_fp64_peak": (C.c_int, [_ctx, _dp]),
    "parsim_pair_peak": (C.c_int, [_ctx, _dp]),
    "parsim_host_init_particles": (C.c_int, [C.c_long, C.c_double, C.c_long, C.c_longlong, C.c_longlong,
                                             C.c_longlong, _dp, _dp, _dp, _dp, _dp]),
    "parsim_create_rows": (C.c_int, [C.POINTER(_ctx), C.c_double, C.c_long, C.c_longlong, C.c_int, C.POINTER(C.c_long)]),
    "parsim_create_rank_rows": (C.c_int, [C.POINTER(_ctx), C.c_double, C.c_long, C.c_longlong, C.c_int, C.c_int,
                                          C.c_int, C.POINTER(C.c_ubyte), C.POINTER(C.c_long)]),
    "parsim_balance_rows": (C.c_int, [C.c_long, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_long)]),
    "parsim_host_row_counts": (C.c_int, [C.c_double, C.c_long, C.c_longlong, _dp, C.POINTER(C.c_longlong)]),
    "parsim_strip_rows": (C.c_int, [C.c_long, C.c_int, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_long)]),
    "parsim_init_device": (C.c_int, [_ctx, C.c_long]),
    "parsim_run": (C.c_int, [_ctx, C.c_long]),
    "parsim_result": (C.c_int, [_ctx, _dp, _dp, C.POINTER(C.c_longlong)]),
    "parsim_debug_stage": (C.c_int, [_ctx, C.c_int]),
    "parsim_dump": (C.c_int, [_ctx, _dp, _dp, _dp, _dp, _dp, _ip, _ip, _ip,
                              C.POINTER(C.c_ulonglong), C.POINTER(C.c_ubyte)]),
    "parsim_dump_cells": (C.c_int, [_ctx, _dp, _dp, _dp]),
    "parsim_kernel_name": (C.c_char_p, [C.c_int]),
    "parsim_stats": (C.c_int, [_ctx, C.POINTER(Stats)]),
    "parsim_stats_reset": (C.c_int, [_ctx]),
    "parsim_set_option": (C.c_int, [_ctx, C.c_char_p, C.c_longlong]),
    "parsim_stream": (C.c_void_p, [_ctx]),
    "parsim_cell_com_static": (C.c_int, [C.c_double, C.c_long, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
}

_lib = None


def load():
    """Load the CUDA library; raises if it was not built (run __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C particlesimulation_b200/csrc` "
                "(or __graft_entry__.build()); particlesimulation_b200 has no CPU fallback")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib
