"""ctypes binding of libkssd_b200.so (the C-ABI in include/kssd_b200.h).  There is no CPU fallback: a missing
library or a missing CUDA device raises."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libkssd_b200.so")

u8p = C.POINTER(C.c_uint8)
u32p = C.POINTER(C.c_uint32)
u64p = C.POINTER(C.c_uint64)
i32p = C.POINTER(C.c_int32)
f64p = C.POINTER(C.c_double)


class Hit(C.Structure):
    _fields_ = [("row", C.c_uint32), ("col", C.c_uint32), ("common", C.c_uint32)]


class Batch(C.Structure):
    _fields_ = [("seq", u8p), ("qual", u8p), ("capacity", C.c_uint64), ("n_bytes", C.c_uint64),
                ("seg_off", u64p), ("seg_genome", u32p), ("seg_capacity", C.c_uint32), ("n_seg", C.c_uint32),
                ("first_emit", C.c_uint64), ("priv", C.c_void_p), ("raw_fasta", C.c_int), ("seg_flags", u8p), ("seg_first", u32p)]


SINK = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(Hit), C.c_size_t)

# name -> (restype, argtypes): every symbol include/kssd_b200.h declares
SYMBOLS = {
    "kssd_b200_last_error": (C.c_char_p, []),
    "kssd_b200_version": (C.c_int, []),
    "kssd_b200_shuffle_table": (C.c_int, [C.c_int, C.c_int, C.c_int, i32p]),
    "kssd_b200_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "kssd_b200_init": (C.c_int, [C.POINTER(C.c_void_p), C.c_int]),
    "kssd_b200_destroy": (None, [C.c_void_p]),
    "kssd_b200_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "kssd_b200_sync": (C.c_int, [C.c_void_p]),
    "kssd_b200_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "kssd_b200_set_force_radix": (C.c_int, [C.c_void_p, C.c_int]),
    "kssd_b200_set_params": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, i32p]),
    "kssd_b200_sketch_begin": (C.c_int, [C.c_void_p, C.c_uint32, C.c_int, C.c_int, C.c_int]),
    "kssd_b200_batch_acquire": (C.c_int, [C.c_void_p, C.POINTER(C.POINTER(Batch)), C.c_uint64, C.c_uint32]),
    "kssd_b200_batch_submit": (C.c_int, [C.c_void_p, C.POINTER(Batch)]),
    "kssd_b200_sketch_scan_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                               C.c_uint32, C.c_uint64]),
    "kssd_b200_sketch_scan_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                             C.c_uint32, C.c_uint64]),
    "kssd_b200_sketch_finish": (C.c_int, [C.c_void_p, C.POINTER(u32p), C.POINTER(u64p)]),
    "kssd_b200_sketch_finish_device": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), u64p]),
    "kssd_b200_sketch_concat_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p),
                                                 C.POINTER(C.c_void_p), u64p, C.POINTER(C.c_int)]),
    "kssd_b200_peer_init": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint64, u8p]),
    "kssd_b200_peer_connect": (C.c_int, [C.c_void_p, u8p, C.c_uint32, C.c_uint32]),
    "kssd_b200_sketch_publish": (C.c_int, [C.c_void_p]),
    "kssd_b200_sketch_gather_peers": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), u64p, C.POINTER(C.c_int)]),
    "kssd_b200_index_build": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int]),
    "kssd_b200_index_build_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32]),
    "kssd_b200_index_build_range": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint64,
                                              u64p, C.POINTER(C.c_int)]),
    "kssd_b200_index_adopt_parts": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32, C.POINTER(C.c_int)]),
    "kssd_b200_index_export": (C.c_int, [C.c_void_p, u64p, u64p, C.POINTER(u32p), C.POINTER(u32p), C.POINTER(u32p)]),
    "kssd_b200_index_load": (C.c_int, [C.c_void_p, C.c_uint64, u32p, u32p, C.c_uint64, u32p, C.c_uint32]),
    "kssd_b200_index_device_views": (C.c_int, [C.c_void_p, u64p, u64p, u32p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                               C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "kssd_b200_index_adopt_device": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p,
                                               C.c_void_p, C.c_void_p]),
    "kssd_b200_set_emit_zero_cells": (C.c_int, [C.c_void_p, C.c_int]),
    "kssd_b200_set_max_neighbor": (C.c_int, [C.c_void_p, C.c_uint64]),
    "kssd_b200_queries_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_void_p)]),
    "kssd_b200_dist_rows": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32,
                                      C.c_int, C.c_int, C.c_double, C.c_int, SINK, C.c_void_p]),
    "kssd_b200_dist_rows_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                             C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_void_p), u64p]),
    "kssd_b200_set_relax_guard": (C.c_int, [C.c_void_p, C.c_int]),
    "kssd_b200_sketch_finish64": (C.c_int, [C.c_void_p, C.POINTER(u64p), C.POINTER(u64p)]),
    "kssd_b200_sketch_finish_device64": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), u64p]),
    "kssd_b200_index_build64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int]),
    "kssd_b200_index_export64": (C.c_int, [C.c_void_p, u64p, u64p, C.POINTER(u64p), C.POINTER(u32p), C.POINTER(u32p)]),
    "kssd_b200_index_load64": (C.c_int, [C.c_void_p, C.c_uint64, u64p, u32p, u32p, C.c_uint64, u32p, C.c_uint32]),
    "kssd_b200_dist_rows64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_uint32, C.c_uint32,
                                        C.c_int, C.c_int, C.c_double, C.c_int, SINK, C.c_void_p]),
    "kssd_b200_dist_rows_device64": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                               C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_void_p), u64p]),
    "kssd_b200_set_union": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), u64p]),
    "kssd_b200_set_subtract": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.c_int,
                                         C.POINTER(C.c_void_p), C.POINTER(u64p)]),
    "kssd_b200_set_subtract_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p, C.c_uint64,
                                                 C.c_void_p, C.c_void_p]),
    "kssd_b200_launch_count": (C.c_uint64, [C.c_void_p]),
    "kssd_b200_scan_kernel_time": (C.c_int, [C.c_void_p, f64p, u64p, u64p]),
    "kssd_b200_count_kernel_time": (C.c_int, [C.c_void_p, f64p, u64p, u64p, u64p]),
}

_lib = None


def load():
    """dlopen the in-tree library and bind every declared symbol.  Raises if the library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                               "(nvcc, sm_100a).  rabbitkssd_b200 has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so does not export it
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


class KssdError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libkssd_b200 error {code}: {msg}")
        self.code = code


def check(rc):
    if rc != 0:
        raise KssdError(rc, load().kssd_b200_last_error().decode(errors="replace"))
