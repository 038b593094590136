"""GroupBy / GroupByRequest with pylibcudf's interface over the C ABI (include/bgb.h).

Mirrors /root/reference/python/pylibcudf/pylibcudf/groupby.pyx:33-214: GroupBy(keys, null_handling).aggregate(requests)
returns (unique-keys Table, [one Table of result columns per request]); group order is arbitrary
(/root/reference/cpp/include/cudf/groupby.hpp:148-150). Errors: LogicError for what libcudf reports as
cudf::logic_error, UnsupportedError where libcudf would take its sort-based fallback.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .aggregation import NullPolicy
from .column import Column, Table, torch_dtype


class GroupByRequest:
    def __init__(self, values: Column, aggregations):
        self._values = values
        self._aggregations = list(aggregations)


class GroupBy:
    def __init__(self, keys: Table, null_handling: int = NullPolicy.EXCLUDE, *, options: dict | None = None):
        self._keys = keys  # keep the keys alive (groupby.pyx:135-137)
        self._null_handling = null_handling
        self._options = options or {}
        self.last_path = None

    def _c_requests(self, requests):
        n = len(requests)
        creq = (_lib.Request * max(n, 1))()
        keep = []
        for i, r in enumerate(requests):
            aggs = (_lib.Aggregation * max(len(r._aggregations), 1))()
            for j, a in enumerate(r._aggregations):
                aggs[j] = _lib.Aggregation(a.kind, a.ddof)
            keep.append(aggs)
            creq[i] = _lib.Request(r._values.view(), aggs, len(r._aggregations), 0)
        return creq, keep

    def aggregate(self, requests, stream=None):
        L = _lib.lib()
        kcols = self._keys.columns()
        nk = len(kcols)
        ckeys = (_lib.ColumnView * max(nk, 1))()
        for i, c in enumerate(kcols):
            ckeys[i] = c.view()
        creq, keep = self._c_requests(requests)
        device = kcols[0].data.device if nk else torch.device("cuda")
        if stream is None:
            stream = torch.cuda.current_stream(device)
        sptr = C.c_void_p(stream.cuda_stream)
        opt = _lib.Options()
        opt.force_path = int(self._options.get("force_path", 0))
        opt.capacity_log2 = int(self._options.get("capacity_log2", 0))
        opt.sample_rows = int(self._options.get("sample_rows", 0))
        handle = C.c_void_p()
        ngroups = C.c_int32(0)
        with torch.cuda.device(device):
            _lib.check(L.bgb_groupby_build(ckeys, nk, int(self._null_handling), creq, len(requests), sptr, None, C.byref(opt),
                                           C.byref(handle), C.byref(ngroups)))
            try:
                G = ngroups.value
                self.last_path = L.bgb_groupby_path(handle).decode()
                with torch.cuda.stream(stream):
                    # caller-allocated outputs (the role of the caller's rmm memory resource)
                    okeys = (_lib.MutableColumn * max(nk, 1))()
                    key_cols = []
                    for i in range(nk):
                        t, sc, nl = C.c_int32(), C.c_int32(), C.c_int32()
                        _lib.check(L.bgb_groupby_key_info(handle, i, C.byref(t), C.byref(sc), C.byref(nl)))
                        data = torch.empty(G, dtype=torch_dtype(t.value), device=device)
                        mask = torch.zeros((G + 31) // 32, dtype=torch.int32, device=device) if nl.value else None
                        key_cols.append(Column(data, mask, 0, t.value, scale=sc.value))
                        okeys[i] = _lib.MutableColumn(t.value, G, data.data_ptr() if G else None,
                                                      mask.data_ptr() if mask is not None and G else None, 0, 0)
                    nres = sum(len(r._aggregations) for r in requests)
                    ores = (_lib.MutableColumn * max(nres, 1))()
                    res_cols = []
                    k = 0
                    for ri, r in enumerate(requests):
                        cols = []
                        for ai in range(len(r._aggregations)):
                            t, sc, nl = C.c_int32(), C.c_int32(), C.c_int32()
                            _lib.check(L.bgb_groupby_result_info(handle, ri, ai, C.byref(t), C.byref(sc), C.byref(nl)))
                            data = torch.empty(G, dtype=torch_dtype(t.value), device=device)
                            mask = torch.zeros((G + 31) // 32, dtype=torch.int32, device=device) if nl.value else None
                            cols.append(Column(data, mask, 0, t.value, scale=sc.value))
                            ores[k] = _lib.MutableColumn(t.value, G, data.data_ptr() if G else None,
                                                         mask.data_ptr() if mask is not None and G else None, 0, 0)
                            k += 1
                        res_cols.append(cols)
                _lib.check(L.bgb_groupby_emit(handle, okeys, ores, sptr))
                for i, c in enumerate(key_cols):
                    c.null_count = okeys[i].null_count
                k = 0
                for cols in res_cols:
                    for c in cols:
                        c.null_count = ores[k].null_count
                        k += 1
            finally:
                L.bgb_groupby_destroy(handle)
        del keep
        return Table(key_cols), [Table(cols) for cols in res_cols]


def can_use_hash_groupby(keys: Table, requests) -> bool:
    L = _lib.lib()
    kcols = keys.columns()
    ckeys = (_lib.ColumnView * max(len(kcols), 1))()
    for i, c in enumerate(kcols):
        ckeys[i] = c.view()
    gb = GroupBy(keys)
    creq, keep = gb._c_requests(requests)
    out = C.c_int32(0)
    _lib.check(L.bgb_can_use_hash_groupby(ckeys, len(kcols), creq, len(requests), C.byref(out)))
    return bool(out.value)
