"""TEST INFRASTRUCTURE ONLY: CPU restatement of cudf::hash_partition's row -> partition assignment.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this module; the
product package (cudf_b200) never does.

What it restates (citations relative to /root/reference):
  * MurmurHash3_x86_32 of an element's bytes       cpp/include/cudf/hashing/detail/murmurhash3_x86_32.cuh:20-110
    (libcudf delegates the byte loop to cuco::murmurhash3_32, a third-party header that is NOT vendored under
    /root/reference - rapids-cmake fetches cuCollections at build time. It is Austin Appleby's public-domain
    MurmurHash3_x86_32: 4-byte little-endian blocks, c1 = 0xcc9e2d51, c2 = 0x1b873593, rotl 15 / 13, h*5+0xe6546b64,
    tail, length xor, fmix32. The restatement is pinned to the published verification vectors of that algorithm in
    tests/test_oracle_hash_partition.py, and to the reference's own documented call sites.)
  * element normalisation: bool -> uint8 0/1, -0.0 -> +0.0, NaN -> quiet NaN
                                                   murmurhash3_x86_32.cuh:46-68, hash_functions.cuh:19-37
  * null element -> UINT32_MAX; first column's hash is the initial value, the others fold in with hash_combine
                                                   cpp/include/cudf/detail/row_operator/hashing.cuh:40-134
  * hash_combine(lhs, rhs) = lhs ^ (rhs + 0x9e3779b9 + (lhs << 6) + (lhs >> 2))
                                                   cpp/include/cudf/hashing/detail/hashing.hpp:83-86
  * partition = hash % num_partitions (bitwise AND when a power of two: same value)
                                                   cpp/src/partitioning/partitioning.cu:41-77
  * IdentityHash: static_cast<uint32_t>(element)   cpp/src/partitioning/partitioning.cu:855-874
  * empty result for no rows / no key columns / num_partitions <= 0
                                                   cpp/src/partitioning/partitioning.cu:885-888
Parity status: the reference cannot be built or run here (libcudf needs cmake + fetched dependencies), and its own tests
for this function are property tests (cpp/tests/partitioning/hash_partition_test.cpp), so the hash is pinned to the
public MurmurHash3 vectors and the partitioning to those properties.
"""
from __future__ import annotations

import numpy as np

M32 = np.uint64(0xFFFFFFFF)


def _rotl(x, r):
    return ((x << np.uint64(r)) | (x >> np.uint64(32 - r))) & M32


def _fmix(h):
    h = h ^ (h >> np.uint64(16))
    h = (h * np.uint64(0x85EBCA6B)) & M32
    h = h ^ (h >> np.uint64(13))
    h = (h * np.uint64(0xC2B2AE35)) & M32
    return h ^ (h >> np.uint64(16))


def murmur3_x86_32_bytes(data: bytes, seed: int = 0) -> int:
    """Scalar MurmurHash3_x86_32 of an arbitrary byte string (the known-answer form)."""
    c1, c2, m = 0xCC9E2D51, 0x1B873593, 0xFFFFFFFF
    h = seed & m
    nblocks = len(data) // 4
    for i in range(nblocks):
        k = int.from_bytes(data[4 * i:4 * i + 4], "little")
        k = (k * c1) & m
        k = ((k << 15) | (k >> 17)) & m
        k = (k * c2) & m
        h ^= k
        h = ((h << 13) | (h >> 19)) & m
        h = (h * 5 + 0xE6546B64) & m
    tail = data[4 * nblocks:]
    k = 0
    if len(tail) >= 3:
        k ^= tail[2] << 16
    if len(tail) >= 2:
        k ^= tail[1] << 8
    if len(tail) >= 1:
        k ^= tail[0]
        k = (k * c1) & m
        k = ((k << 15) | (k >> 17)) & m
        k = (k * c2) & m
        h ^= k
    h ^= len(data)
    h ^= h >> 16
    h = (h * 0x85EBCA6B) & m
    h ^= h >> 13
    h = (h * 0xC2B2AE35) & m
    h ^= h >> 16
    return h


def murmur3_x86_32_fixed(elements: np.ndarray, seed: int = 0) -> np.ndarray:
    """Vectorised MurmurHash3_x86_32 of every element of a fixed-width array (1, 2, 4 or 8 bytes each) -> uint32."""
    a = np.ascontiguousarray(elements)
    width = a.dtype.itemsize
    raw = a.view({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[width]).astype(np.uint64)
    c1, c2 = np.uint64(0xCC9E2D51), np.uint64(0x1B873593)
    h = np.full(raw.shape, seed & 0xFFFFFFFF, dtype=np.uint64)

    def mixk(k):
        k = (k * c1) & M32
        k = _rotl(k, 15)
        return (k * c2) & M32

    if width >= 4:
        blocks = [raw & M32] + ([raw >> np.uint64(32)] if width == 8 else [])
        for k in blocks:
            h = h ^ mixk(k)
            h = _rotl(h, 13)
            h = (h * np.uint64(5) + np.uint64(0xE6546B64)) & M32
    else:
        h = h ^ mixk(raw)
    h = h ^ np.uint64(width)
    return _fmix(h).astype(np.uint32)


def normalize(values: np.ndarray) -> np.ndarray:
    """bool -> uint8 0/1, -0.0 -> +0.0, every NaN -> the quiet NaN (murmurhash3_x86_32.cuh:46-68)."""
    v = np.asarray(values)
    if v.dtype == np.bool_:
        return v.astype(np.uint8)
    if v.dtype.kind == "f":
        out = v.copy()
        out[v == 0] = 0.0
        out[np.isnan(v)] = np.float32(np.nan) if v.dtype == np.float32 else np.nan
        bits = out.view(np.uint32 if v.dtype == np.float32 else np.uint64).copy()
        bits[np.isnan(v)] = 0x7FC00000 if v.dtype == np.float32 else 0x7FF8000000000000
        return bits
    return v


def identity_hash(values: np.ndarray) -> np.ndarray:
    """static_cast<uint32_t>(element) for integers and bool (sign-extended, truncated to 32 bits)."""
    v = np.asarray(values)
    if v.dtype == np.bool_:
        return v.astype(np.uint32)
    if v.dtype.kind not in "iu":
        raise TypeError("identity hash restated for integer / bool elements only (float -> uint32 casts are UB off-range)")
    return v.astype(np.int64 if v.dtype.kind == "i" else np.uint64).view(np.uint64).astype(np.uint64).__and__(M32).astype(np.uint32)


def hash_combine(lhs: np.ndarray, rhs: np.ndarray) -> np.ndarray:
    l, r = lhs.astype(np.uint64), rhs.astype(np.uint64)
    return (l ^ ((r + np.uint64(0x9E3779B9) + ((l << np.uint64(6)) & M32) + (l >> np.uint64(2))) & M32)).astype(np.uint32)


def row_hash(columns, seed: int = 0, hash_function: str = "murmur3") -> np.ndarray:
    """columns: list of (values ndarray, valid bool ndarray | None) in hashing order -> uint32 row hashes."""
    h = None
    for values, valid in columns:
        e = murmur3_x86_32_fixed(normalize(values), seed) if hash_function == "murmur3" else identity_hash(values)
        if valid is not None:
            e = np.where(np.asarray(valid, dtype=bool), e, np.uint32(0xFFFFFFFF)).astype(np.uint32)
        h = e if h is None else hash_combine(h, e)
    return h


def partition_numbers(columns, num_partitions: int, seed: int = 0, hash_function: str = "murmur3") -> np.ndarray:
    return (row_hash(columns, seed, hash_function).astype(np.uint64) % np.uint64(num_partitions)).astype(np.int64)


def hash_partition(table, columns_to_hash, num_partitions: int, seed: int = 0, hash_function: str = "murmur3"):
    """table: list of (values, valid | None). Returns (row order, offsets[num_partitions + 1]): a stable ordering of the
    input rows by partition number (the reference leaves the order inside a partition unspecified)."""
    n = len(table[0][0]) if table else 0
    if num_partitions <= 0 or n == 0 or not columns_to_hash:
        return np.zeros(0, dtype=np.int64), [0] * (max(num_partitions, 0) + 1)
    pid = partition_numbers([table[i] for i in columns_to_hash], num_partitions, seed, hash_function)
    order = np.argsort(pid, kind="stable")
    counts = np.bincount(pid, minlength=num_partitions)
    return order, [0] + np.cumsum(counts).tolist()
