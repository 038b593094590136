"""Device columns and tables with the cudf layout (data buffer + 32-bit-word validity bitmask).

Mirrors the parts of pylibcudf.Column / pylibcudf.Table the groupby path needs
(/root/reference/python/pylibcudf/pylibcudf/column.pyx, table.pyx); torch tensors are only the device buffers.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib


class TypeId:
    """cudf::type_id (/root/reference/cpp/include/cudf/types.hpp:185-217)."""
    EMPTY = 0
    INT8, INT16, INT32, INT64 = 1, 2, 3, 4
    UINT8, UINT16, UINT32, UINT64 = 5, 6, 7, 8
    FLOAT32, FLOAT64, BOOL8 = 9, 10, 11
    TIMESTAMP_DAYS, TIMESTAMP_SECONDS, TIMESTAMP_MILLISECONDS, TIMESTAMP_MICROSECONDS, TIMESTAMP_NANOSECONDS = 12, 13, 14, 15, 16
    DURATION_DAYS, DURATION_SECONDS, DURATION_MILLISECONDS, DURATION_MICROSECONDS, DURATION_NANOSECONDS = 17, 18, 19, 20, 21
    DICTIONARY32, STRING, LIST, DECIMAL32, DECIMAL64, DECIMAL128, STRUCT = 22, 23, 24, 25, 26, 27, 28


_TORCH_OF_TYPE = {
    1: torch.int8, 2: torch.int16, 3: torch.int32, 4: torch.int64, 5: torch.uint8, 6: torch.uint16, 7: torch.uint32,
    8: torch.uint64, 9: torch.float32, 10: torch.float64, 11: torch.bool, 12: torch.int32, 13: torch.int64, 14: torch.int64,
    15: torch.int64, 16: torch.int64, 17: torch.int32, 18: torch.int64, 19: torch.int64, 20: torch.int64, 21: torch.int64,
    25: torch.int32, 26: torch.int64,
}
_TYPE_OF_TORCH = {torch.int8: 1, torch.int16: 2, torch.int32: 3, torch.int64: 4, torch.uint8: 5, torch.uint16: 6,
                  torch.uint32: 7, torch.uint64: 8, torch.float32: 9, torch.float64: 10, torch.bool: 11}
_NP_OF_TYPE = {1: np.int8, 2: np.int16, 3: np.int32, 4: np.int64, 5: np.uint8, 6: np.uint16, 7: np.uint32, 8: np.uint64,
               9: np.float32, 10: np.float64, 11: np.bool_, 12: np.int32, 13: np.int64, 14: np.int64, 15: np.int64,
               16: np.int64, 17: np.int32, 18: np.int64, 19: np.int64, 20: np.int64, 21: np.int64, 25: np.int32, 26: np.int64}


def torch_dtype(type_id: int):
    return _TORCH_OF_TYPE[type_id]


def pack_mask(valid: np.ndarray) -> np.ndarray:
    """bool[n] -> bitmask words (int32 view of uint32), bit i%32 of word i//32, 1 = valid (cudf/utilities/bit.hpp:47-60)."""
    n = len(valid)
    words = (n + 31) // 32
    padded = np.zeros(max(words, 1) * 32, dtype=np.uint8)
    padded[:n] = np.asarray(valid, dtype=np.uint8)
    return np.packbits(padded.reshape(-1, 8), axis=1, bitorder="little").reshape(-1).view(np.int32).copy()


def unpack_mask(words: np.ndarray, n: int) -> np.ndarray:
    bits = np.unpackbits(words.view(np.uint8), bitorder="little")
    return bits[:n].astype(bool)


class Column:
    """Non-owning-view semantics of cudf::column_view over torch device buffers."""

    def __init__(self, data: torch.Tensor, mask: torch.Tensor | None = None, null_count: int = 0, type_id: int | None = None,
                 offset: int = 0, size: int | None = None, scale: int = 0):
        self.data = data
        self.mask = mask  # int32 words on the same device
        self.type_id = type_id if type_id is not None else _TYPE_OF_TORCH[data.dtype]
        self.offset = offset
        self.size = (data.numel() - offset) if size is None else size
        self.null_count = null_count if mask is not None else 0
        self.scale = scale
        self.children = []  # STRUCT columns (SUM_OVERFLOW results): [sum Column, overflow Column]; `data` is then empty

    # -- construction helpers (tests, examples)
    @staticmethod
    def from_numpy(arr: np.ndarray, valid: np.ndarray | None = None, device="cuda", type_id: int | None = None) -> "Column":
        a = np.ascontiguousarray(arr)
        if a.dtype == np.uint16 or a.dtype == np.uint32 or a.dtype == np.uint64:
            t = torch.from_numpy(a.view({2: np.int16, 4: np.int32, 8: np.int64}[a.dtype.itemsize])).to(device)
            t = t.view({2: torch.uint16, 4: torch.uint32, 8: torch.uint64}[a.dtype.itemsize])
        else:
            t = torch.from_numpy(a).to(device)
        if a.size == 0:
            t = torch.empty(0, dtype=t.dtype, device=device)
        mask, nc = None, 0
        if valid is not None:
            v = np.asarray(valid, dtype=bool)
            mask = torch.from_numpy(pack_mask(v)).to(device)
            nc = int(len(v) - np.count_nonzero(v))
        return Column(t, mask, nc, type_id)

    def sliced(self, offset: int, size: int) -> "Column":
        """cudf::slice: same buffers, new offset/size (null_count recomputed by the caller when it matters)."""
        c = Column(self.data, self.mask, self.null_count, self.type_id, self.offset + offset, size, self.scale)
        return c

    def to_numpy(self):
        """-> (values ndarray, valid bool ndarray)"""
        n = self.size
        t = self.data[self.offset:self.offset + n]
        if t.dtype in (torch.uint16, torch.uint32, torch.uint64):
            host = t.view({torch.uint16: torch.int16, torch.uint32: torch.int32, torch.uint64: torch.int64}[t.dtype]).cpu().numpy()
            host = host.view({2: np.uint16, 4: np.uint32, 8: np.uint64}[host.dtype.itemsize])
        else:
            host = t.cpu().numpy()
        if self.mask is None:
            valid = np.ones(n, dtype=bool)
        else:
            bits = np.unpackbits(self.mask.cpu().numpy().view(np.uint8), bitorder="little")
            valid = bits[self.offset:self.offset + n].astype(bool)
        return host, valid

    @staticmethod
    def dictionary(indices: "Column", keys: "Column", mask: torch.Tensor | None = None, null_count: int = 0) -> "Column":
        """cudf dictionary column (cudf/dictionary/dictionary_column_view.hpp): row r holds keys[indices[r]]; the parent carries the
        null mask. Served as a VALUE column of a groupby request; results are typed by `keys`."""
        c = Column(torch.empty(0, dtype=torch.int32, device=indices.data.device), mask, null_count, TypeId.DICTIONARY32, 0, indices.size)
        c.children = [indices, keys]
        return c

    def view(self) -> _lib.ColumnView:
        if self.type_id == TypeId.DICTIONARY32:
            # `data` = host array of the two child views, kept alive by this Column
            self._child_views = (_lib.ColumnView * 2)(self.children[0].view(), self.children[1].view())
            return _lib.ColumnView(self.type_id, self.scale, self.size, self.null_count, self.offset, 0,
                                   C.cast(self._child_views, C.c_void_p).value, self.mask.data_ptr() if self.mask is not None else None)
        return _lib.ColumnView(self.type_id, self.scale, self.size, self.null_count, self.offset, 0,
                               self.data.data_ptr() if self.data.numel() else None,
                               self.mask.data_ptr() if self.mask is not None else None)


class Table:
    def __init__(self, columns):
        self._columns = list(columns)

    def columns(self):
        return self._columns

    def num_columns(self):
        return len(self._columns)

    def num_rows(self):
        return self._columns[0].size if self._columns else 0
