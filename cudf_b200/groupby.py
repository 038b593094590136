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
from .column import Column, Table, TypeId, torch_dtype


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
        self.last_profile = []  # [(kernel name, device ms)] when options["profile"] is set
        self._peer = None       # distributed.PeerContext: build through bgb_groupby_build_peer (collective)

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
        opt.profile = int(self._options.get("profile", 0))
        opt.bucket_slice_log2 = int(self._options.get("bucket_slice_log2", 0))
        handle = C.c_void_p()
        ngroups = C.c_int32(0)
        with torch.cuda.device(device):
            if self._peer is not None:
                _lib.check(L.bgb_groupby_build_peer(self._peer.handle, ckeys, nk, int(self._null_handling), creq, len(requests), sptr,
                                                    None, C.byref(opt), C.byref(handle), C.byref(ngroups)))
            else:
                _lib.check(L.bgb_groupby_build(ckeys, nk, int(self._null_handling), creq, len(requests), sptr, None, C.byref(opt),
                                               C.byref(handle), C.byref(ngroups)))
            try:
                self.last_path = L.bgb_groupby_path(handle).decode()
                key_cols, res_cols = emit_handle(L, handle, ngroups.value, nk, [len(r._aggregations) for r in requests], device, stream,
                                                 struct_child_types=[r._values.type_id for r in requests])
                if opt.profile:
                    self.last_profile = _read_profile(L, handle)
            finally:
                L.bgb_groupby_destroy(handle)
        del keep
        return Table(key_cols), [Table(cols) for cols in res_cols]


def emit_handle(L, handle, G, nk, aggs_per_request, device, stream, struct_child_types=None):
    """Caller-allocated outputs (the role of the caller's rmm memory resource) + bgb_groupby_emit for a built handle.
    Returns ([key Column], [[result Column] per request])."""
    sptr = C.c_void_p(stream.cuda_stream)
    keep_alive = []
    with torch.cuda.stream(stream):
        okeys = (_lib.MutableColumn * max(nk, 1))()
        key_cols = []
        for i in range(nk):
            t, sc, nl = C.c_int32(), C.c_int32(), C.c_int32()
            _lib.check(L.bgb_groupby_key_info(handle, i, C.byref(t), C.byref(sc), C.byref(nl)))
            data = torch.empty(G, dtype=torch_dtype(t.value), device=device)
            mask = torch.empty((G + 31) // 32, dtype=torch.int32, device=device) if nl.value else None  # emit writes every word
            key_cols.append(Column(data, mask, 0, t.value, scale=sc.value))
            okeys[i] = _lib.MutableColumn(t.value, G, data.data_ptr() if G else None,
                                          mask.data_ptr() if mask is not None and G else None, 0, 0)
        nres = sum(aggs_per_request)
        ores = (_lib.MutableColumn * max(nres, 1))()
        res_cols = []
        k = 0
        for ri, na in enumerate(aggs_per_request):
            cols = []
            for ai in range(na):
                t, sc, nl = C.c_int32(), C.c_int32(), C.c_int32()
                _lib.check(L.bgb_groupby_result_info(handle, ri, ai, C.byref(t), C.byref(sc), C.byref(nl)))
                mask = torch.empty((G + 31) // 32, dtype=torch.int32, device=device) if nl.value else None  # emit writes every word
                if t.value == TypeId.STRUCT:
                    # SUM_OVERFLOW: struct{sum: value type, overflow: BOOL8}; `data` = the two child descriptors (bgb.h)
                    vt = struct_child_types[ri] if struct_child_types else None
                    kids = [torch.empty(G, dtype=torch_dtype(vt), device=device), torch.empty(G, dtype=torch.bool, device=device)]
                    cdesc = (_lib.MutableColumn * 2)(_lib.MutableColumn(vt, G, kids[0].data_ptr() if G else None, None, 0, 0),
                                                     _lib.MutableColumn(TypeId.BOOL8, G, kids[1].data_ptr() if G else None, None, 0, 0))
                    keep_alive.append(cdesc)
                    col = Column(torch.empty(0, dtype=torch.int8, device=device), mask, 0, t.value, size=G)
                    col.children = [Column(kids[0], None, 0, vt, scale=sc.value), Column(kids[1], None, 0, TypeId.BOOL8)]
                    cols.append(col)
                    ores[k] = _lib.MutableColumn(t.value, G, C.cast(cdesc, C.c_void_p), mask.data_ptr() if mask is not None and G else None, 0, 0)
                    k += 1
                    continue
                data = torch.empty(G, dtype=torch_dtype(t.value), device=device)
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
    return key_cols, res_cols


def _read_profile(L, handle):
    out, i = [], 0
    name, ms = C.c_char_p(), C.c_float()
    while L.bgb_groupby_profile(handle, i, C.byref(name), C.byref(ms)) == 0:
        out.append((name.value.decode(), ms.value))
        i += 1
    return out


class HostColumnView:
    """A host-memory column (numpy data + optional packed uint32 validity words) for groupby_host."""

    def __init__(self, data, mask_words=None, null_count=0, type_id=None):
        from .column import _NP_OF_TYPE
        self.data = data
        self.mask = mask_words
        self.null_count = null_count if mask_words is not None else 0
        if type_id is None:
            type_id = {v: k for k, v in reversed(list(_NP_OF_TYPE.items()))}[data.dtype.type]
        self.type_id = type_id

    def view(self):
        return _lib.ColumnView(self.type_id, 0, len(self.data), self.null_count, 0, 0, self.data.ctypes.data if len(self.data) else None,
                               self.mask.ctypes.data if self.mask is not None else None)


def groupby_host(keys, requests, null_handling: int = NullPolicy.EXCLUDE, options: dict | None = None):
    """End-to-end call on HOST columns through the C ABI's bgb_groupby_host: H2D staging, build, emit, D2H.

    keys: [HostColumnView]; requests: [(HostColumnView, [Aggregation])].
    Returns (num_groups, [(data, mask_words|None, null_count)] keys, [[...] per request]) as numpy arrays that VIEW the host
    buffers the library allocated (no copy); the buffers are released when the last array is garbage-collected."""
    import numpy as np
    from .column import _NP_OF_TYPE
    L = _lib.lib()
    options = options or {}
    nk = len(keys)
    ckeys = (_lib.ColumnView * max(nk, 1))()
    for i, c in enumerate(keys):
        ckeys[i] = c.view()
    creq = (_lib.Request * max(len(requests), 1))()
    keep = []
    for i, (col, aggs) in enumerate(requests):
        ca = (_lib.Aggregation * max(len(aggs), 1))()
        for j, a in enumerate(aggs):
            ca[j] = _lib.Aggregation(a.kind, a.ddof)
        keep.append(ca)
        creq[i] = _lib.Request(col.view(), ca, len(aggs), 0)
    opt = _lib.Options()
    opt.force_path = int(options.get("force_path", 0))
    opt.capacity_log2 = int(options.get("capacity_log2", 0))
    opt.bucket_slice_log2 = int(options.get("bucket_slice_log2", 0))
    ng, nres = C.c_int32(0), C.c_int32(0)
    hk, hr = C.POINTER(_lib.HostColumn)(), C.POINTER(_lib.HostColumn)()
    _lib.check(L.bgb_groupby_host(ckeys, nk, int(null_handling), creq, len(requests), C.byref(opt), C.byref(ng), C.byref(hk),
                                  C.byref(hr), C.byref(nres)))
    G = ng.value

    class _Owner:
        """keeps the C result alive while any returned array views it; frees it (bgb_free_host_result) afterwards"""
        def __init__(self):
            self.hk, self.hr, self.nk, self.nres = hk, hr, nk, nres.value
        def __del__(self):
            L.bgb_free_host_result(self.hk, self.nk)
            L.bgb_free_host_result(self.hr, self.nres)

    owner = _Owner()

    def view(ptr, nbytes, dt):
        buf = (C.c_uint8 * nbytes).from_address(ptr)
        buf._owner = owner  # the numpy array references `buf` (buffer protocol), `buf` references the owner
        return np.frombuffer(buf, dtype=dt)

    def take(hc):
        dt = np.dtype(_NP_OF_TYPE[hc.type_id])
        if G == 0:
            return np.empty(0, dt), None, 0
        data = view(hc.data, G * dt.itemsize, dt)   # zero-copy: the library's host buffer
        mask = None
        if hc.null_mask:
            mask = view(hc.null_mask, ((G + 31) // 32) * 4, np.uint32)
        return data, mask, hc.null_count

    kout = [take(hk[i]) for i in range(nk)]
    rout, k = [], 0
    for _, aggs in requests:
        rout.append([take(hr[k + j]) for j in range(len(aggs))])
        k += len(aggs)
    del owner
    return G, kout, rout


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
