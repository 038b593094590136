"""`hash_partition` with pylibcudf's signature and the reference's partition assignment, bit for bit.

Mirrors pylibcudf.partitioning.hash_partition (/root/reference/python/pylibcudf/pylibcudf/partitioning.pyx:33-102) over
`bgb_hash_partition` (include/bgb.h), which replaces cudf::hash_partition
(/root/reference/cpp/include/cudf/partitioning.hpp:103-145, cpp/src/partitioning/partitioning.cu:567-760,876-950).
There is no CPU fallback: without libbgb.so or a GPU the call raises.
"""
from __future__ import annotations

import ctypes as C
import enum

import torch

from . import _lib
from .column import Column, Table

DEFAULT_HASH_SEED = 0  # cudf::DEFAULT_HASH_SEED (/root/reference/cpp/include/cudf/hashing.hpp:29)


class HashId(enum.IntEnum):
    """cudf::hash_id (/root/reference/cpp/include/cudf/partitioning.hpp:28-31)."""
    HASH_IDENTITY = 0
    HASH_MURMUR3 = 1


def hash_partition(input: Table, keys, num_partitions: int, hash_function: HashId = HashId.HASH_MURMUR3,
                   seed: int = DEFAULT_HASH_SEED, stream=None):
    """Partition the rows of `input` by the hash of `keys`.

    keys: list of column indices of `input`, or a Table of key columns with as many rows as `input`.
    Returns (Table, offsets) with num_partitions + 1 offsets; partition i is rows [offsets[i], offsets[i + 1]).
    Zero rows, zero key columns or num_partitions <= 0 give an empty table and all-zero offsets
    (partitioning.cu:885-888). The order of rows inside a partition is unspecified, as in the reference.
    """
    L = _lib.lib()
    cols = list(input.columns())
    if isinstance(keys, Table):
        kcols = keys.columns()
        if kcols and kcols[0].size != input.num_rows():
            raise ValueError("Input table and key table must have same number of rows, or key table should have no columns.")
        # a key table that is not made of input's own columns rides along as extra columns and is dropped afterwards
        index = []
        for k in kcols:
            at = next((i for i, c in enumerate(cols) if c is k), None)
            if at is None:
                cols.append(k)
                at = len(cols) - 1
            index.append(at)
    else:
        index = [int(i) for i in keys]
        for i in index:
            if i < 0 or i >= len(cols):
                raise IndexError(f"columns_to_hash index {i} is out of range for a table of {len(cols)} columns")
    n_out_cols = input.num_columns()
    nparts = int(num_partitions)
    n = input.num_rows()
    offsets = (C.c_int32 * (max(nparts, 0) + 1))()
    if nparts <= 0 or n == 0 or not index:
        empty = [Column(torch.empty(0, dtype=c.data.dtype, device=c.data.device), None, 0, c.type_id, scale=c.scale) for c in cols[:n_out_cols]]
        return Table(empty), [0] * (max(nparts, 0) + 1)
    device = cols[0].data.device
    if device.type != "cuda":
        raise RuntimeError("hash_partition runs on a CUDA device only (there is no CPU path)")
    cv = (_lib.ColumnView * len(cols))()
    oc = (_lib.MutableColumn * len(cols))()
    outs = []
    for i, c in enumerate(cols):
        cv[i] = c.view()
        data = torch.empty(n, dtype=c.data.dtype, device=device)
        mask = torch.empty((n + 31) // 32, dtype=torch.int32, device=device) if c.mask is not None and c.null_count != 0 else None
        outs.append((data, mask))
        oc[i] = _lib.MutableColumn(c.type_id, n, data.data_ptr(), mask.data_ptr() if mask is not None else None, 0, 0)
    idx = (C.c_int32 * len(index))(*index)
    with torch.cuda.device(device):
        sp = stream.cuda_stream if stream is not None else torch.cuda.current_stream(device).cuda_stream
        _lib.check(L.bgb_hash_partition(cv, len(cols), idx, len(index), nparts, int(hash_function), int(seed) & 0xFFFFFFFF, oc,
                                        offsets, C.c_void_p(sp)))
    result = [Column(d, m, oc[i].null_count if m is not None else 0, cols[i].type_id, scale=cols[i].scale)
              for i, (d, m) in enumerate(outs[:n_out_cols])]
    return Table(result), [int(offsets[i]) for i in range(nparts + 1)]
