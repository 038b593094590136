"""CPU-side checks of the drop-in boundary: libbgb.so loads and exports every entry point include/bgb.h declares
(no compute calls: there is no GPU here), and the host-side planner rules match the reference's type table."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "bgb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bgb_[a-z0-9_]+)\s*\(", src)) - {"bgb_allocator"})


def test_library_exports_every_declared_symbol():
    from cudf_b200 import _lib
    L = _lib.lib()
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(L, n), f"include/bgb.h declares {n} but libbgb.so does not export it"


def test_version_and_error_string_without_gpu():
    from cudf_b200 import _lib
    L = _lib.lib()
    assert L.bgb_version() == 100
    L.bgb_last_error.restype = C.c_char_p
    assert isinstance(L.bgb_last_error(), bytes)


@pytest.mark.parametrize("tid,kind,expect", [
    (3, 0, 4),    # SUM(int32) -> int64            aggregation.hpp:878-1012
    (1, 0, 4),    # SUM(int8)  -> int64
    (10, 0, 10),  # SUM(float64) -> float64
    (9, 0, 9),    # SUM(float32) -> float32
    (3, 5, 3),    # COUNT_VALID -> int32 (size_type)
    (10, 6, 3),   # COUNT_ALL -> int32
    (2, 3, 2),    # MIN keeps the source type
    (4, 10, 10),  # MEAN(int64) -> float64
    (3, 12, 10),  # VARIANCE -> float64
    (3, 13, 10),  # STD -> float64
    (13, 0, -1),  # SUM(timestamp) invalid
    (23, 12, -1),  # VARIANCE(string) invalid
    (18, 10, 18),  # MEAN(duration) keeps duration
])
def test_target_type_rules_match_oracle_and_reference(tid, kind, expect):
    from cudf_b200 import _lib
    from oracle import oracle as orc
    L = _lib.lib()
    assert L.bgb_target_type(tid, kind) == expect
    if tid != 23:
        assert orc.lib().orc_target_type(tid, kind) == expect


@pytest.mark.parametrize("tid,expect", [(1, 28), (2, 28), (3, 28), (4, 28), (25, 28), (26, 28), (7, -1), (11, -1), (10, -1), (13, -1)])
def test_sum_overflow_target_type_and_eligibility(tid, expect):
    # SUM_OVERFLOW -> STRUCT{sum, overflow} for signed integers and decimal32/64 only (sum_overflow_tests.cpp:496-561: bool and
    # unsigned types are rejected); the hash path says so without touching a GPU
    from cudf_b200 import _lib
    L = _lib.lib()
    assert L.bgb_target_type(tid, 1) == expect
    keys = (_lib.ColumnView * 1)(_lib.ColumnView(3, 0, 4, 0, 0, 0, None, None))
    aggs = (_lib.Aggregation * 1)(_lib.Aggregation(1, 1))
    reqs = (_lib.Request * 1)(_lib.Request(_lib.ColumnView(tid, 0, 4, 0, 0, 0, None, None), aggs, 1, 0))
    ok = C.c_int32(-1)
    rc = L.bgb_can_use_hash_groupby(keys, 1, reqs, 1, C.byref(ok))
    if expect == 28:
        assert rc == 0 and ok.value == 1
    else:
        assert ok.value == 0 or rc != 0
