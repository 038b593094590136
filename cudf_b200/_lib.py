"""ctypes binding of the C ABI in include/bgb.h (libbgb.so, built in-tree by __graft_entry__.build()).

There is NO fallback: if the CUDA library is missing, importing the operators raises.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbgb.so")


class BgbError(RuntimeError):
    """Maps the C status codes onto the reference's exception classes (cudf/utilities/error.hpp:35-97)."""

    def __init__(self, status: int, message: str):
        super().__init__(message)
        self.status = status


class LogicError(BgbError):      # cudf::logic_error
    pass


class UnsupportedError(BgbError):  # can_use_hash_groupby == false
    pass


class CudaError(BgbError):       # cudf::cuda_error
    pass


class PeerOverflow(BgbError):    # BGB_PEER_OVERFLOW: every rank of the exchange reports it in the same step
    pass


class PeerAborted(CudaError):    # BGB_PEER_ABORTED: another rank failed in this step (every other rank raises this together)
    pass


class ColumnView(C.Structure):
    _fields_ = [("type_id", C.c_int32), ("scale", C.c_int32), ("size", C.c_int32), ("null_count", C.c_int32),
                ("offset", C.c_int32), ("reserved", C.c_int32), ("data", C.c_void_p), ("null_mask", C.c_void_p)]


class Aggregation(C.Structure):
    _fields_ = [("kind", C.c_int32), ("ddof", C.c_int32)]


class Request(C.Structure):
    _fields_ = [("values", ColumnView), ("aggregations", C.POINTER(Aggregation)), ("num_aggregations", C.c_int32),
                ("reserved", C.c_int32)]


class MutableColumn(C.Structure):
    _fields_ = [("type_id", C.c_int32), ("size", C.c_int32), ("data", C.c_void_p), ("null_mask", C.c_void_p),
                ("null_count", C.c_int32), ("reserved", C.c_int32)]


class StreamRequest(C.Structure):
    _fields_ = [("column_index", C.c_int32), ("kind", C.c_int32), ("ddof", C.c_int32), ("reserved", C.c_int32)]


class Options(C.Structure):
    _fields_ = [("force_path", C.c_int32), ("capacity_log2", C.c_int32), ("sample_rows", C.c_int32), ("profile", C.c_int32),
                ("bucket_slice_log2", C.c_int32), ("reserved", C.c_int32 * 3)]


class HostColumn(C.Structure):
    _fields_ = [("type_id", C.c_int32), ("size", C.c_int32), ("data", C.c_void_p), ("null_mask", C.c_void_p),
                ("null_count", C.c_int32), ("reserved", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.bgb_last_error.restype = C.c_char_p
    L.bgb_groupby_path.restype = C.c_char_p
    L.bgb_groupby_path.argtypes = [C.c_void_p]
    L.bgb_groupby_build.argtypes = [C.POINTER(ColumnView), C.c_int32, C.c_int32, C.POINTER(Request), C.c_int32, C.c_void_p,
                                    C.c_void_p, C.POINTER(Options), C.POINTER(C.c_void_p), C.POINTER(C.c_int32)]
    L.bgb_groupby_result_info.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32),
                                          C.POINTER(C.c_int32)]
    L.bgb_groupby_key_info.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.bgb_groupby_emit.argtypes = [C.c_void_p, C.POINTER(MutableColumn), C.POINTER(MutableColumn), C.c_void_p]
    L.bgb_groupby_destroy.argtypes = [C.c_void_p]
    L.bgb_groupby_destroy.restype = None
    L.bgb_can_use_hash_groupby.argtypes = [C.POINTER(ColumnView), C.c_int32, C.POINTER(Request), C.c_int32, C.POINTER(C.c_int32)]
    L.bgb_target_type.argtypes = [C.c_int32, C.c_int32]
    L.bgb_target_type.restype = C.c_int32
    L.bgb_gen_uniform_int.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int64, C.c_void_p]
    L.bgb_gen_uniform_real.argtypes = [C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_double, C.c_void_p]
    L.bgb_gen_null_mask.argtypes = [C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint32, C.POINTER(C.c_int32), C.c_void_p]
    L.bgb_gen_categorical_i8.argtypes = [C.c_void_p, C.c_int64, C.c_uint64, C.c_uint64, C.POINTER(C.c_int32), C.c_char_p,
                                         C.c_int32, C.c_void_p]
    L.bgb_groupby_host.argtypes = [C.POINTER(ColumnView), C.c_int32, C.c_int32, C.POINTER(Request), C.c_int32,
                                   C.POINTER(Options), C.POINTER(C.c_int32), C.POINTER(C.POINTER(HostColumn)),
                                   C.POINTER(C.POINTER(HostColumn)), C.POINTER(C.c_int32)]
    L.bgb_free_host_result.argtypes = [C.POINTER(HostColumn), C.c_int32]
    L.bgb_free_host_result.restype = None
    L.bgb_groupby_profile.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_float)]
    L.bgb_kernel_launches.restype = C.c_int64
    L.bgb_set_estimate_speculation.argtypes = [C.c_int32]
    L.bgb_set_estimate_speculation.restype = C.c_int32
    L.bgb_partition_by_key_hash.argtypes = [C.POINTER(ColumnView), C.c_int32, C.POINTER(ColumnView), C.c_int32, C.c_int32,
                                            C.POINTER(MutableColumn), C.POINTER(C.c_int32), C.c_void_p]
    L.bgb_hash_partition.argtypes = [C.POINTER(ColumnView), C.c_int32, C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32, C.c_uint32,
                                     C.POINTER(MutableColumn), C.POINTER(C.c_int32), C.c_void_p]
    L.bgb_bytes_to_bitmask.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_int32), C.c_void_p]
    L.bgb_bytes_to_bitmask_noalloc.argtypes = L.bgb_bytes_to_bitmask.argtypes
    L.bgb_bytes_to_bitmask_noalloc_max_rows.restype = C.c_int64
    L.bgb_finalize_moments.argtypes = [C.c_int32, C.c_int32, C.POINTER(ColumnView), C.POINTER(ColumnView), C.POINTER(ColumnView),
                                       C.POINTER(MutableColumn), C.c_void_p]
    L.bgb_is_streaming_groupby_supported.argtypes = [C.c_int32, C.c_int32, C.POINTER(C.c_int32)]
    L.bgb_stream_create.argtypes = [C.POINTER(C.c_int32), C.c_int32, C.POINTER(StreamRequest), C.c_int32, C.c_int32, C.c_int32,
                                    C.POINTER(C.c_void_p)]
    L.bgb_stream_aggregate.argtypes = [C.c_void_p, C.POINTER(ColumnView), C.c_int32, C.c_void_p, C.POINTER(Options)]
    L.bgb_stream_merge.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.bgb_stream_distinct_keys.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
    L.bgb_stream_finalize.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32)]
    L.bgb_stream_destroy.argtypes = [C.c_void_p]
    L.bgb_stream_destroy.restype = None
    L.bgb_peer_create.argtypes = [C.c_int32, C.c_int32, C.c_int64, C.POINTER(C.c_void_p)]
    L.bgb_peer_export.argtypes = [C.c_void_p, C.c_void_p]
    L.bgb_peer_attach.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.bgb_peer_attach_local.argtypes = [C.c_void_p, C.c_int32, C.c_void_p]
    L.bgb_peer_segment_capacity.argtypes = [C.c_void_p, C.c_int32]
    L.bgb_peer_segment_capacity.restype = C.c_int64
    L.bgb_peer_destroy.argtypes = [C.c_void_p]
    L.bgb_peer_destroy.restype = None
    L.bgb_groupby_build_peer.argtypes = [C.c_void_p] + list(L.bgb_groupby_build.argtypes)
    _lib = L
    return L


def check(status: int):
    if status == 0:
        return
    msg = lib().bgb_last_error().decode()
    if status == 1:
        raise LogicError(status, msg)
    if status == 2:
        raise UnsupportedError(status, msg)
    if status == 6:
        raise PeerOverflow(status, msg)
    if status == 7:
        raise PeerAborted(status, msg)
    raise CudaError(status, msg)
