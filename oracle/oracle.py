"""TEST INFRASTRUCTURE ONLY: ctypes front end of the CPU oracle (oracle/groupby_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module. The product package (cudf_b200) never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# cudf::type_id (cpp/include/cudf/types.hpp:185-217)
TYPE_ID = {
    "int8": 1, "int16": 2, "int32": 3, "int64": 4, "uint8": 5, "uint16": 6, "uint32": 7, "uint64": 8,
    "float32": 9, "float64": 10, "bool": 11,
}
NP_OF_TYPE_ID = {
    1: np.int8, 2: np.int16, 3: np.int32, 4: np.int64, 5: np.uint8, 6: np.uint16, 7: np.uint32, 8: np.uint64,
    9: np.float32, 10: np.float64, 11: np.bool_, 12: np.int32, 13: np.int64, 14: np.int64, 15: np.int64, 16: np.int64,
    17: np.int32, 18: np.int64, 19: np.int64, 20: np.int64, 21: np.int64, 25: np.int32, 26: np.int64,
}
# cudf::aggregation::Kind (cpp/include/cudf/aggregation.hpp:78-121)
KIND = {"sum": 0, "product": 2, "min": 3, "max": 4, "count_valid": 5, "count_all": 6, "sum_of_squares": 9,
        "mean": 10, "m2": 11, "variance": 12, "std": 13, "argmax": 16, "argmin": 17,
        # SUM_OVERFLOW yields struct{sum, overflow}: "sum_overflow" is the sum child (typed like the values), "sum_overflow_flag" the
        # BOOL8 overflow child; both carry the struct's validity
        "sum_overflow": 1, "sum_overflow_flag": 101}


class _Col(C.Structure):
    _fields_ = [("type_id", C.c_int32), ("size", C.c_int32), ("data", C.c_void_p), ("null_mask", C.c_void_p),
                ("null_count", C.c_int32), ("offset", C.c_int32)]


class _Agg(C.Structure):
    _fields_ = [("value_index", C.c_int32), ("kind", C.c_int32), ("ddof", C.c_int32)]


class _Out(C.Structure):
    _fields_ = [("type_id", C.c_int32), ("has_mask", C.c_int32), ("data", C.c_void_p), ("valid", C.c_void_p)]


class _Res(C.Structure):
    _fields_ = [("num_groups", C.c_int32), ("n_keys", C.c_int32), ("n_aggs", C.c_int32),
                ("keys", C.POINTER(_Out)), ("aggs", C.POINTER(_Out))]


def build() -> str:
    """Compile the oracle shared library in place (gcc only; no reference sources involved)."""
    so = os.path.join(_HERE, "libgroupby_oracle.so")
    src = os.path.join(_HERE, "groupby_oracle.c")
    hdr = os.path.join(_HERE, "groupby_oracle.h")
    if (not os.path.exists(so)) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.orc_groupby.restype = C.c_int
        _LIB.orc_groupby.argtypes = [C.POINTER(_Col), C.c_int, C.c_int, C.POINTER(_Col), C.c_int, C.POINTER(_Agg),
                                     C.c_int, C.c_int, C.c_int, C.POINTER(_Res)]
        _LIB.orc_free_result.argtypes = [C.POINTER(_Res)]
        _LIB.orc_last_error.restype = C.c_char_p
        _LIB.orc_splitmix64.restype = C.c_uint64
        _LIB.orc_splitmix64.argtypes = [C.c_uint64, C.c_uint64]
        _LIB.orc_target_type.restype = C.c_int
        _LIB.orc_target_type.argtypes = [C.c_int, C.c_int]
    return _LIB


def pack_mask(valid: np.ndarray) -> np.ndarray:
    """bool[n] -> cudf bitmask words (uint32, bit i of word i//32, 1 = valid; cudf/utilities/bit.hpp:47-60)."""
    n = len(valid)
    words = (n + 31) // 32
    padded = np.zeros(words * 32, dtype=np.uint8)
    padded[:n] = np.asarray(valid, dtype=np.uint8)
    return np.packbits(padded.reshape(-1, 8), axis=1, bitorder="little").reshape(-1).view(np.uint32).copy()


@dataclass
class HostColumn:
    """Host column with cudf layout: data + optional bitmask."""
    data: np.ndarray
    valid: np.ndarray | None = None  # bool per row, None = no null mask
    type_id: int | None = None

    def tid(self) -> int:
        if self.type_id is not None:
            return self.type_id
        return TYPE_ID[str(self.data.dtype)]


class OracleError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(msg)
        self.code = code


def groupby(keys, values, aggs, include_null_keys=False, threads=1, sort=True):
    """keys/values: lists of HostColumn; aggs: list of (value_index, kind_name, ddof).

    Returns (num_groups, key_cols, agg_cols) where each col is (np.ndarray data, np.ndarray bool valid, has_mask).
    Rows are key-sorted (nulls first) when sort=True.
    """
    L = lib()
    keep = []

    def mk(cols):
        arr = (_Col * max(len(cols), 1))()
        for i, c in enumerate(cols):
            d = np.ascontiguousarray(c.data)
            if d.dtype == np.bool_:
                d = d.view(np.uint8)
            keep.append(d)
            m = None
            nc = 0
            if c.valid is not None:
                m = pack_mask(c.valid)
                nc = int(len(c.valid) - np.count_nonzero(c.valid))
                keep.append(m)
            arr[i] = _Col(c.tid(), len(d), d.ctypes.data, m.ctypes.data if m is not None else None, nc, 0)
        return arr

    ck, cv = mk(keys), mk(values)
    ca = (_Agg * max(len(aggs), 1))()
    for i, (vi, kind, ddof) in enumerate(aggs):
        ca[i] = _Agg(vi, KIND[kind] if isinstance(kind, str) else kind, ddof)
    res = _Res()
    rc = L.orc_groupby(ck, len(keys), int(include_null_keys), cv, len(values), ca, len(aggs), threads, int(sort),
                       C.byref(res))
    if rc != 0:
        raise OracleError(rc, L.orc_last_error().decode())
    G = res.num_groups

    def take(o):
        dt = NP_OF_TYPE_ID[o.type_id]
        if dt is np.bool_:
            data = np.ctypeslib.as_array(C.cast(o.data, C.POINTER(C.c_uint8)), shape=(max(G, 1),))[:G].astype(np.bool_)
        else:
            nbytes = np.dtype(dt).itemsize * max(G, 1)
            raw = np.ctypeslib.as_array(C.cast(o.data, C.POINTER(C.c_uint8)), shape=(nbytes,))
            data = raw.view(dt)[:G].copy()
        valid = np.ctypeslib.as_array(C.cast(o.valid, C.POINTER(C.c_uint8)), shape=(max(G, 1),))[:G].astype(np.bool_)
        return data, valid, bool(o.has_mask)

    kout = [take(res.keys[i]) for i in range(res.n_keys)]
    aout = [take(res.aggs[i]) for i in range(res.n_aggs)]
    L.orc_free_result(C.byref(res))
    return G, kout, aout


def splitmix64(seed: int, rows: np.ndarray) -> np.ndarray:
    """Vectorised counter-based SplitMix64 x(seed,row); identical to orc_splitmix64 and the GPU generators."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) * np.uint64(0x1000003) + rows.astype(np.uint64)
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))
