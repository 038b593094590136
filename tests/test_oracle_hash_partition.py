"""CPU: pins oracle/hash_partition.py (the checker of the GPU hash_partition) to golden vectors.

1. Published MurmurHash3_x86_32 verification vectors (the algorithm libcudf gets from cuco::murmurhash3_32).
2. The reference's own golden row hashes: MurmurHashTest.BasicList / NullableList
   (/root/reference/cpp/tests/hashing/murmurhash3_x86_32_test.cpp:212-300). A LIST<UINT64> row hashes as
   combine(combine(0, murmur(int32 size, seed 0)), murmur(uint64 element, seed))... with the same element hasher,
   seed handling and hash_combine the fixed-width path uses (row_operator/hashing.cuh:179-212), so these values pin the
   4-byte and 8-byte murmur, the seed and the combine step against output of the reference itself.
3. The reference's known-answer partitions: python/cudf/cudf/tests/dataframe/methods/test_partition_by_hash.py:41-133.
"""
import numpy as np
import pytest

from oracle import hash_partition as hp

PUBLIC_VECTORS = [
    (b"", 0, 0x00000000), (b"", 1, 0x514E28B7), (b"", 0xFFFFFFFF, 0x81F16F39),
    (b"\x00\x00\x00\x00", 0, 0x2362F9DE), (b"\xff\xff\xff\xff", 0, 0x76293B50),
    (b"\x21\x43\x65\x87", 0, 0xF55B516B), (b"\x21\x43\x65\x87", 0x5082EDEE, 0x2362F9DE),
    (b"\x21\x43\x65", 0, 0x7E4A8634), (b"\x21\x43", 0, 0xA0F7B07A), (b"\x21", 0, 0x72661CF4),
    (b"\x00\x00\x00", 0, 0x85F0B427), (b"\x00\x00", 0, 0x30F4C306), (b"\x00", 0, 0x514E28B7),
    (b"Hello, world!", 1234, 0xFAF6CDB3), (b"Hello, world!", 4321, 0xBF505788),
    (b"The quick brown fox jumps over the lazy dog", 0x9747B28C, 0x2FA826CD),
    (b"aaaa", 0x9747B28C, 0x5A97808A), (b"abc", 0, 0xB3DD93FA),
    (b"abcdbcdecdefdefgefghfghighijhijkijkljklmklmnlmnomnopnopq", 0, 0xEE925B90),
]


@pytest.mark.parametrize("data,seed,expect", PUBLIC_VECTORS)
def test_murmur3_public_vectors(data, seed, expect):
    assert hp.murmur3_x86_32_bytes(data, seed) == expect


@pytest.mark.parametrize("dtype", [np.uint8, np.int16, np.uint32, np.int64, np.float32, np.float64])
@pytest.mark.parametrize("seed", [0, 7, 0xDEADBEEF])
def test_vectorised_matches_scalar(dtype, seed):
    rng = np.random.default_rng(3)
    a = rng.integers(0, 255, 64 * np.dtype(dtype).itemsize, dtype=np.uint8).view(dtype)
    got = hp.murmur3_x86_32_fixed(a, seed)
    want = np.array([hp.murmur3_x86_32_bytes(x.tobytes(), seed) for x in a], dtype=np.uint32)
    assert (got == want).all()


def _list_row_hash(elements, seed, valid=None):
    """row hash of one LIST<UINT64> row (row_operator/hashing.cuh:179-212); valid=None: the column has no nulls"""
    h = np.array([0], dtype=np.uint32)
    if valid is not None:
        h = hp.hash_combine(h, np.array([0 if valid else 0xFFFFFFFF], dtype=np.uint32))
    h = hp.hash_combine(h, hp.murmur3_x86_32_fixed(np.array([len(elements)], dtype=np.int32), 0))
    for e in elements:
        h = hp.hash_combine(h, hp.murmur3_x86_32_fixed(np.array([e], dtype=np.uint64), seed))
    return int(h[0])


def test_reference_golden_basic_list():
    rows = [[], [], [1], [1, 1], [1], [1, 2], [2, 2], [2], [2], [2, 1], [2, 2], [2, 2]]
    expect = [3248124823, 3248124823, 1004521430, 1508072170, 1004521430, 3926647866, 616744158, 2401729030, 2401729030,
              2980709038, 616744158, 616744158]
    expect_seeded = [3248124823, 3248124823, 3217320641, 2844201578, 3217320641, 3748008630, 3397385561, 3869469325,
                     3869469325, 836697381, 3397385561, 3397385561]
    assert [_list_row_hash(r, 0) for r in rows] == expect
    assert [_list_row_hash(r, 15) for r in rows] == expect_seeded


def test_reference_golden_nullable_list():
    rows = [[], [], [1], [1], [2, 2], [2], [2], [], [2, 2], [2, 2], []]
    valids = [True, True, True, True, True, True, True, False, True, True, False]
    expect = [3912350204, 3912350204, 1608859631, 1608859631, 435283192, 3506305375, 3506305375, 3912350141, 435283192,
              435283192, 3912350141]
    expect_seeded = [3912350204, 3912350204, 2678848247, 2678848247, 2235667558, 373403129, 373403129, 3912350141,
                     2235667558, 2235667558, 3912350141]
    assert [_list_row_hash(r, 0, v) for r, v in zip(rows, valids)] == expect
    assert [_list_row_hash(r, 31, v) for r, v in zip(rows, valids)] == expect_seeded


def _parts(table, cols, n):
    order, offsets = hp.hash_partition(table, cols, n)
    return [order[offsets[i]:offsets[i + 1]].tolist() for i in range(n)], offsets


def test_reference_kat_masked_value():
    # test_dataframe_hash_partition_masked_value: key = arange(10), 3 partitions -> {9}, {0, 8}, {1..7}
    parts, offsets = _parts([(np.arange(10, dtype=np.int64), None)], [0], 3)
    assert parts == [[9], [0, 8], [1, 2, 3, 4, 5, 6, 7]]
    assert offsets == [0, 1, 3, 10]


def test_reference_kat_masked_keys():
    # test_dataframe_hash_partition_masked_keys: keys [0, 1, 2, null, null] -> {null, null}, {0}, {1, 2}
    rng = np.random.default_rng(seed=0)
    null = rng.choice([True, False], size=5)
    assert null.tolist() == [False, False, False, True, True]  # what the expected frames of the reference test encode
    parts, _ = _parts([(np.arange(5, dtype=np.int64), ~null)], [0], 3)
    assert parts == [[3, 4], [0], [1, 2]]


def test_reference_kat_keep_index():
    # test_dataframe_hash_partition_keep_index: key [3, 2, 1, 4, 5], 2 partitions -> keys {2, 1, 4} and {3, 5}
    parts, _ = _parts([(np.array([3, 2, 1, 4, 5], dtype=np.int64), None)], [0], 2)
    assert parts == [[1, 2, 3], [0, 4]]


def test_normalisation_and_nulls():
    for dt in (np.float32, np.float64):
        a = np.array([0.0, np.nan, 1.5], dtype=dt)
        b = np.array([-0.0, -np.nan, 1.5], dtype=dt)
        assert (hp.row_hash([(a, None)]) == hp.row_hash([(b, None)])).all()
    t = np.array([True, False])
    assert (hp.row_hash([(t, None)]) == hp.murmur3_x86_32_fixed(np.array([1, 0], dtype=np.uint8))).all()
    v = np.arange(4, dtype=np.int32)
    h = hp.row_hash([(v, np.array([True, False, True, False]))], seed=5)
    assert h[1] == 0xFFFFFFFF and h[3] == 0xFFFFFFFF and h[0] == hp.murmur3_x86_32_fixed(v[:1], 5)[0]


def test_empty_results():
    t = [(np.arange(4, dtype=np.int32), None)]
    assert hp.hash_partition(t, [0], 0)[1] == [0]
    assert hp.hash_partition(t, [], 3)[1] == [0, 0, 0, 0]
    assert hp.hash_partition([(np.zeros(0, dtype=np.int32), None)], [0], 2)[1] == [0, 0, 0]


def test_identity_hash():
    assert hp.identity_hash(np.array([-1, 5], dtype=np.int8)).tolist() == [0xFFFFFFFF, 5]
    assert hp.identity_hash(np.array([2**40 + 3], dtype=np.int64)).tolist() == [3]
    assert hp.partition_numbers([(np.arange(10, dtype=np.int32), None)], 4, hash_function="identity").tolist() == [i % 4 for i in range(10)]
