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


def test_eligibility_of_wide_keys_and_dictionary_values_without_gpu():
    """bgb_can_use_hash_groupby (hash/groupby.cu:115-131) only plans: four int64 key columns (256 bits of key: the row-number table)
    and a DICTIONARY32 value column (typed by its keys, hash/groupby.cu:118-120) are served; a dictionary of strings, a dictionary
    without its child views and a ninth key column are not."""
    from cudf_b200 import _lib
    L = _lib.lib()
    INT64, UINT8, FLOAT64, DICT, STRING = 4, 5, 10, 22, 23
    SUM, MIN = 0, 2

    def can(keys, values, kind):
        aggs = (_lib.Aggregation * 1)(_lib.Aggregation(kind, 1))
        reqs = (_lib.Request * 1)(_lib.Request(values, aggs, 1, 0))
        ok = C.c_int32(-1)
        rc = L.bgb_can_use_hash_groupby(keys, len(keys), reqs, 1, C.byref(ok))
        return rc, ok.value

    plain = _lib.ColumnView(INT64, 0, 4, 0, 0, 0, None, None)
    four = (_lib.ColumnView * 4)(*[_lib.ColumnView(INT64, 0, 4, 0, 0, 0, None, None) for _ in range(4)])
    assert can(four, plain, SUM) == (0, 1)
    nine = (_lib.ColumnView * 9)(*[_lib.ColumnView(INT64, 0, 4, 0, 0, 0, None, None) for _ in range(9)])
    assert can(nine, plain, SUM)[1] == 0
    one = (_lib.ColumnView * 1)(plain)
    children = (_lib.ColumnView * 2)(_lib.ColumnView(UINT8, 0, 4, 0, 0, 0, None, None), _lib.ColumnView(FLOAT64, 0, 3, 0, 0, 0, None, None))
    dcol = _lib.ColumnView(DICT, 0, 4, 0, 0, 0, C.cast(children, C.c_void_p).value, None)
    assert can(one, dcol, SUM) == (0, 1) and can(one, dcol, MIN) == (0, 1)
    str_children = (_lib.ColumnView * 2)(_lib.ColumnView(UINT8, 0, 4, 0, 0, 0, None, None), _lib.ColumnView(STRING, 0, 3, 0, 0, 0, None, None))
    sdict = _lib.ColumnView(DICT, 0, 4, 0, 0, 0, C.cast(str_children, C.c_void_p).value, None)
    rc, ok = can(one, sdict, MIN)
    assert ok == 0 or rc != 0
    rc, ok = can(one, _lib.ColumnView(DICT, 0, 4, 0, 0, 0, None, None), SUM)
    assert ok == 0 or rc != 0
