"""GPU tests of StreamingGroupBy (bgb_stream_*), transcribed from the reference's own streaming tests
(/root/reference/cpp/tests/groupby/streaming_groupby_test.cpp) plus oracle comparisons on random multi-batch inputs:
like the reference's verify_against_groupby (:60-110), the expected answer is the stateless groupby of the concatenated
batches -- here computed by the CPU oracle."""
import numpy as np
import pytest

import cudf_b200 as cb
from cudf_b200 import aggregation as agg
from oracle import oracle as orc
from tests.gpu_util import make_agg, sort_order

pytestmark = pytest.mark.gpu

MAXK = 1024


def col(a, valid=None, dtype=None):
    return cb.Column.from_numpy(np.asarray(a, dtype=dtype), None if valid is None else np.asarray(valid, dtype=bool))


def table(*cols):
    return cb.Table(list(cols))


def single(kind, ddof=1, column=1):
    return [cb.StreamingAggregationRequest(column, make_agg(kind, ddof))]


def finalize_sorted(sg):
    kt, rts = sg.finalize()
    keys = [c.to_numpy() for c in kt.columns()]
    order = sort_order(keys)
    return [(d[order], v[order]) for d, v in keys], [tuple(x[order] for x in rt.columns()[0].to_numpy()) for rt in rts]


def verify_against_groupby(sg, batches, key_cols, reqs, include_null_keys=False, rtol=1e-9):
    """batches: list of lists of (data, valid|None) per column."""
    ncols = len(batches[0])
    cat = []
    for c in range(ncols):
        data = np.concatenate([b[c][0] for b in batches])
        if any(b[c][1] is not None for b in batches):
            valid = np.concatenate([b[c][1] if b[c][1] is not None else np.ones(len(b[c][0]), bool) for b in batches])
        else:
            valid = None
        cat.append(orc.HostColumn(data, valid))
    vcols = sorted({r.column_index for r in reqs})
    aggs = []
    for r in reqs:
        aggs.append((vcols.index(r.column_index), r.aggregation.kind, r.aggregation.ddof))
    Go, ko, ao = orc.groupby([cat[k] for k in key_cols], [cat[v] for v in vcols], aggs, include_null_keys=include_null_keys, threads=2)
    keys, res = finalize_sorted(sg)
    assert len(keys[0][0]) == Go
    for c in range(len(key_cols)):
        np.testing.assert_array_equal(keys[c][1], ko[c][1])
        np.testing.assert_array_equal(keys[c][0][keys[c][1]], ko[c][0][ko[c][1]])
    for i in range(len(reqs)):
        np.testing.assert_array_equal(res[i][1], ao[i][1], err_msg=f"request {i} validity")
        got, exp = res[i][0][res[i][1]], ao[i][0][ao[i][1]]
        if got.dtype.kind == "f":
            np.testing.assert_allclose(got, exp, rtol=rtol, atol=0)
        else:
            np.testing.assert_array_equal(got, exp)
    assert sg.distinct_keys() == Go


def feed(sg, batches):
    for b in batches:
        sg.aggregate(table(*[cb.Column.from_numpy(d, v) for d, v in b]))


def b(*cols):
    return [(np.asarray(c[0]), None if c[1] is None else np.asarray(c[1], bool)) if isinstance(c, tuple) else (np.asarray(c), None) for c in cols]


I32, F64 = np.int32, np.float64


@pytest.mark.parametrize("kind", ["sum", "min", "max", "count_valid", "count_all", "mean", "product", "variance", "std", "m2", "sum_of_squares"])
def test_two_batches_every_kind(kind):  # SumTwoBatches, MinMaxTwoBatches, CountValidTwoBatches, MeanTwoBatches, ProductTwoBatches, VarianceBasic ...
    vdt = F64 if kind in ("mean", "variance", "std", "m2", "product") else I32
    batches = [b(np.array([1, 2, 3, 1], I32), np.array([10, 20, 30, 40], vdt)), b(np.array([2, 3, 1, 4], I32), np.array([5, 15, 25, 35], vdt))]
    reqs = single(kind)
    sg = cb.StreamingGroupBy([0], reqs, MAXK)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)


def test_merge_two_objects_known_answer():  # MergeTwoObjects: keys {1,2,3} sums {40,60,50}
    w1 = cb.StreamingGroupBy([0], single("sum"), MAXK)
    w1.aggregate(table(col([1, 2, 1], dtype=I32), col([10, 20, 30], dtype=I32)))
    w2 = cb.StreamingGroupBy([0], single("sum"), MAXK)
    w2.aggregate(table(col([2, 3], dtype=I32), col([40, 50], dtype=I32)))
    w1.merge(w2)
    keys, res = finalize_sorted(w1)
    np.testing.assert_array_equal(keys[0][0], [1, 2, 3])
    np.testing.assert_array_equal(res[0][0], [40, 60, 50])
    assert res[0][0].dtype == np.int64
    # the other object is not modified
    k2, r2 = finalize_sorted(w2)
    np.testing.assert_array_equal(k2[0][0], [2, 3])
    np.testing.assert_array_equal(r2[0][0], [40, 50])


@pytest.mark.parametrize("kind", ["mean", "count_valid", "variance"])
def test_merge_matches_single_object(kind):  # MergeMeanTwoBatches, MergeCountTwoBatches, MergeVarianceTwoBatches
    rng = np.random.default_rng(3)
    batches = [b(rng.integers(0, 20, 500).astype(I32), (rng.standard_normal(500), rng.random(500) > 0.2)) for _ in range(4)]
    reqs = single(kind)
    w1, w2 = cb.StreamingGroupBy([0], reqs, MAXK), cb.StreamingGroupBy([0], reqs, MAXK)
    feed(w1, batches[:2])
    feed(w2, batches[2:])
    w1.merge(w2)
    verify_against_groupby(w1, batches, [0], reqs)


def test_empty_batch_and_single_batch():  # EmptyBatch, SingleBatch
    sg = cb.StreamingGroupBy([0], single("sum"), MAXK)
    batches = [b(np.array([1, 2, 1], I32), np.array([1, 2, 3], I32))]
    sg.aggregate(table(col([], dtype=I32), col([], dtype=I32)))
    feed(sg, batches)
    sg.aggregate(table(col([], dtype=I32), col([], dtype=I32)))
    verify_against_groupby(sg, batches, [0], single("sum"))


def test_new_keys_in_later_batches_and_disjoint_batches():  # NewKeysInLaterBatches, DisjointKeysAcrossBatches, AllDuplicateKeysAcrossBatches
    batches = [b(np.array([1, 2], I32), np.array([1, 2], I32)), b(np.array([3, 4], I32), np.array([3, 4], I32)),
               b(np.array([1, 4, 5], I32), np.array([10, 40, 50], I32)), b(np.array([5, 5, 5], I32), np.array([1, 1, 1], I32))]
    reqs = single("sum")
    sg = cb.StreamingGroupBy([0], reqs, MAXK)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)
    assert sg.distinct_keys() == 5


def test_multiple_requests_on_different_columns_and_same_column():  # MultipleRequestsOnDifferentColumns, SumAndMeanOnSameColumn
    rng = np.random.default_rng(5)
    batches = [b(rng.integers(0, 6, 300).astype(I32), rng.integers(0, 100, 300).astype(np.int64), rng.standard_normal(300)) for _ in range(3)]
    reqs = [cb.StreamingAggregationRequest(1, agg.sum()), cb.StreamingAggregationRequest(2, agg.max()), cb.StreamingAggregationRequest(1, agg.mean()),
            cb.StreamingAggregationRequest(2, agg.std(0)), cb.StreamingAggregationRequest(1, agg.sum())]
    sg = cb.StreamingGroupBy([0], reqs, MAXK)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)


def test_finalize_does_not_modify_state():  # FinalizeDoesNotModifyState
    batches = [b(np.array([1, 2, 1], I32), np.array([10, 20, 30], I32)), b(np.array([2, 3], I32), np.array([40, 50], I32))]
    reqs = single("sum")
    sg = cb.StreamingGroupBy([0], reqs, MAXK)
    feed(sg, batches[:1])
    verify_against_groupby(sg, batches[:1], [0], reqs)
    verify_against_groupby(sg, batches[:1], [0], reqs)  # twice: same answer
    feed(sg, batches[1:])
    verify_against_groupby(sg, batches, [0], reqs)


def test_null_keys_excluded_included_and_all_null():  # NullKeysExcluded, NullKeysIncluded, AllNullKeysExcluded
    batches = [b((np.array([1, 2, 3], I32), [True, False, True]), np.array([10, 20, 30], I32)), b((np.array([1, 2], I32), [True, False]), np.array([40, 50], I32))]
    reqs = single("sum")
    sg = cb.StreamingGroupBy([0], reqs, MAXK)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)
    inc = cb.StreamingGroupBy([0], reqs, MAXK, cb.NullPolicy.INCLUDE)
    feed(inc, batches)
    keys, res = finalize_sorted(inc)  # known answer of the reference test: keys {null, 1, 3}, sums {70, 50, 30} (nulls first here)
    np.testing.assert_array_equal(keys[0][1], [False, True, True])
    np.testing.assert_array_equal(keys[0][0][1:], [1, 3])
    np.testing.assert_array_equal(res[0][0], [70, 50, 30])
    verify_against_groupby(inc, batches, [0], reqs, include_null_keys=True)
    alln = cb.StreamingGroupBy([0], reqs, MAXK)
    alln.aggregate(table(col([1, 2], [False, False], I32), col([1, 2], dtype=I32)))
    keys, res = finalize_sorted(alln)
    assert len(keys[0][0]) == 0 and alln.distinct_keys() == 0


def test_mask_only_in_a_later_batch():
    """The table layout must not depend on which batch happens to carry a null mask."""
    rng = np.random.default_rng(9)
    batches = [b(rng.integers(0, 5, 100).astype(I32), rng.standard_normal(100)),
               b((rng.integers(0, 5, 100).astype(I32), rng.random(100) > 0.1), (rng.standard_normal(100), rng.random(100) > 0.3))]
    for include in (False, True):
        reqs = [cb.StreamingAggregationRequest(1, agg.sum()), cb.StreamingAggregationRequest(1, agg.mean()), cb.StreamingAggregationRequest(1, agg.count())]
        sg = cb.StreamingGroupBy([0], reqs, MAXK, cb.NullPolicy.INCLUDE if include else cb.NullPolicy.EXCLUDE)
        feed(sg, batches)
        verify_against_groupby(sg, batches, [0], reqs, include_null_keys=include)


@pytest.mark.parametrize("vtype", ["int8", "int16", "int32", "int64", "uint8", "uint32", "uint64", "float32", "float64"])
def test_typed_sum_and_min(vtype):  # StreamingGroupbySumTypedTest, StreamingGroupbyMinTypedTest
    batches = [b(np.array([1, 2, 3, 1], I32), np.array([1, 2, 3, 4]).astype(vtype)), b(np.array([2, 3, 1, 4], I32), np.array([5, 6, 7, 8]).astype(vtype))]
    for kind in ("sum", "min"):
        reqs = single(kind)
        sg = cb.StreamingGroupBy([0], reqs, MAXK)
        feed(sg, batches)
        verify_against_groupby(sg, batches, [0], reqs)


def test_unsupported_agg_and_limits_raise():  # UnsupportedAggThrows, BatchExceedsMaxDistinctKeysThrows, ExceedsDistinctKeyCapacityThrows, FinalizeBeforeAggregateThrows
    with pytest.raises(cb.LogicError):
        cb.StreamingGroupBy([0], [cb.StreamingAggregationRequest(1, agg.argmax())], MAXK)
    with pytest.raises(cb.LogicError):
        cb.StreamingGroupBy([0], single("sum"), 0)
    sg = cb.StreamingGroupBy([0], single("sum"), 4)
    with pytest.raises(cb.LogicError):
        sg.finalize()
    with pytest.raises(cb.LogicError):
        sg.aggregate(table(col([0, 1, 2, 3, 4], dtype=I32), col([1, 1, 1, 1, 1], dtype=I32)))  # 5 rows > max_distinct_keys
    sg.aggregate(table(col([0, 1, 2, 3], dtype=I32), col([10, 20, 30, 40], dtype=I32)))
    assert sg.distinct_keys() == 4
    sg.aggregate(table(col([3, 3, 0], dtype=I32), col([1, 1, 1], dtype=I32)))  # no new key: fine
    with pytest.raises(cb.LogicError):
        sg.aggregate(table(col([4], dtype=I32), col([50], dtype=I32)))         # 4 + 1 > 4
    assert cb.is_streaming_groupby_supported(cb.TypeId.INT32, agg.Kind.SUM)
    assert not cb.is_streaming_groupby_supported(cb.TypeId.INT32, agg.Kind.ARGMAX)
    assert not cb.is_streaming_groupby_supported(cb.TypeId.STRING, agg.Kind.SUM)


def test_cumulative_rows_can_exceed_max_distinct_keys_many_small_batches():  # CumulativeRowsCanExceedMaxDistinctKeys, ManySmallBatches, SingleRowBatches
    rng = np.random.default_rng(12)
    batches = [b(rng.integers(0, 8, 8).astype(I32), rng.integers(0, 50, 8).astype(I32)) for _ in range(40)]
    batches += [b(np.array([k], I32), np.array([k * 3], I32)) for k in range(8)]
    reqs = single("sum")
    sg = cb.StreamingGroupBy([0], reqs, 8)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)


def test_sliced_input_columns():  # SlicedInputColumns
    k = cb.Column.from_numpy(np.array([9, 9, 1, 2, 1, 9], I32)).sliced(2, 3)
    v = cb.Column.from_numpy(np.array([7, 7, 10, 20, 30, 7], I32)).sliced(2, 3)
    sg = cb.StreamingGroupBy([0], single("sum"), MAXK)
    sg.aggregate(table(k, v))
    keys, res = finalize_sorted(sg)
    np.testing.assert_array_equal(keys[0][0], [1, 2])
    np.testing.assert_array_equal(res[0][0], [40, 20])


def test_multi_column_keys_and_value_nulls():  # MultiColumnKeys, StdWithNullValues
    rng = np.random.default_rng(15)
    batches = [b(rng.integers(0, 4, 400).astype(I32), rng.integers(0, 3, 400).astype(np.int8), (rng.standard_normal(400), rng.random(400) > 0.25)) for _ in range(3)]
    reqs = [cb.StreamingAggregationRequest(2, agg.std(1)), cb.StreamingAggregationRequest(2, agg.sum()), cb.StreamingAggregationRequest(2, agg.count())]
    sg = cb.StreamingGroupBy([0, 1], reqs, MAXK)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0, 1], reqs)


def test_large_batches_take_the_fast_kernels():
    """1e6-row batches: each batch goes through the shared-memory / tiny kernels before it is folded into the state."""
    rng = np.random.default_rng(21)
    batches = [b(rng.integers(0, 5000, 1_000_000).astype(np.int64), rng.integers(-100, 100, 1_000_000).astype(np.int64)) for _ in range(3)]
    reqs = [cb.StreamingAggregationRequest(1, agg.sum()), cb.StreamingAggregationRequest(1, agg.count())]
    sg = cb.StreamingGroupBy([0], reqs, 1 << 20)
    feed(sg, batches)
    verify_against_groupby(sg, batches, [0], reqs)
