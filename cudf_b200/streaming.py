"""StreamingGroupBy: the reference's stateful streaming groupby over the C ABI (bgb_stream_*).

Mirrors cudf::groupby::streaming_groupby (/root/reference/cpp/include/cudf/groupby.hpp:428-598): constructed from key
column INDICES and one (column index, aggregation) pair per request, fed with tables batch by batch, merged with other
objects' partial state, finalised (without modifying the state) into (unique keys Table, [one result Table per request]).
Errors follow the reference: unsupported kinds, an oversized batch and too many distinct keys raise LogicError
(std::invalid_argument / cudf::logic_error there).
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from .aggregation import NullPolicy
from .column import Table
from .groupby import emit_handle


class StreamingAggregationRequest:
    """streaming_aggregation_request (groupby.hpp:419-422)."""

    def __init__(self, column_index: int, aggregation):
        self.column_index = int(column_index)
        self.aggregation = aggregation


def is_streaming_groupby_supported(type_id: int, kind: int) -> bool:
    out = C.c_int32(0)
    _lib.check(_lib.lib().bgb_is_streaming_groupby_supported(int(type_id), int(kind), C.byref(out)))
    return bool(out.value)


class StreamingGroupBy:
    def __init__(self, key_indices, requests, max_distinct_keys: int, null_handling: int = NullPolicy.EXCLUDE, *, options: dict | None = None):
        L = _lib.lib()
        self._nk = len(key_indices)
        self._nreq = len(requests)
        self._options = options or {}
        ki = (C.c_int32 * max(self._nk, 1))(*[int(k) for k in key_indices])
        rq = (_lib.StreamRequest * max(self._nreq, 1))()
        for i, r in enumerate(requests):
            rq[i] = _lib.StreamRequest(r.column_index, r.aggregation.kind, r.aggregation.ddof, 0)
        self._h = C.c_void_p()
        _lib.check(L.bgb_stream_create(ki, self._nk, rq, self._nreq, int(max_distinct_keys), int(null_handling), C.byref(self._h)))
        self._device = None

    def close(self):
        if getattr(self, "_h", None):
            _lib.lib().bgb_stream_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def aggregate(self, data: Table, stream=None):
        L = _lib.lib()
        cols = data.columns()
        cv = (_lib.ColumnView * max(len(cols), 1))()
        for i, c in enumerate(cols):
            cv[i] = c.view()
        device = cols[0].data.device
        self._device = device
        if stream is None:
            stream = torch.cuda.current_stream(device)
        opt = _lib.Options()
        opt.force_path = int(self._options.get("force_path", 0))
        with torch.cuda.device(device):
            _lib.check(L.bgb_stream_aggregate(self._h, cv, len(cols), C.c_void_p(stream.cuda_stream), C.byref(opt)))

    def merge(self, other: "StreamingGroupBy", stream=None):
        device = self._device or other._device or torch.device("cuda")
        if stream is None:
            stream = torch.cuda.current_stream(device)
        with torch.cuda.device(device):
            _lib.check(_lib.lib().bgb_stream_merge(self._h, other._h, C.c_void_p(stream.cuda_stream)))

    def distinct_keys(self) -> int:
        out = C.c_int32(0)
        _lib.check(_lib.lib().bgb_stream_distinct_keys(self._h, C.byref(out)))
        return out.value

    def finalize(self, stream=None):
        L = _lib.lib()
        device = self._device or torch.device("cuda")
        if stream is None:
            stream = torch.cuda.current_stream(device)
        handle, ng = C.c_void_p(), C.c_int32(0)
        with torch.cuda.device(device):
            _lib.check(L.bgb_stream_finalize(self._h, C.c_void_p(stream.cuda_stream), C.byref(handle), C.byref(ng)))
            try:
                key_cols, res_cols = emit_handle(L, handle, ng.value, self._nk, [1] * self._nreq, device, stream)
            finally:
                L.bgb_groupby_destroy(handle)
        return Table(key_cols), [Table(c) for c in res_cols]
