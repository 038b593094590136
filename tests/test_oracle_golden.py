"""Pins the CPU oracle against the reference's own known-answer tests."""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import kat

CASES = kat.load_cases()


@pytest.mark.parametrize("name,c,kt,vt", CASES, ids=[x[0] for x in CASES])
def test_oracle_matches_reference_kat(name, c, kt, vt):
    keys, vals, exp = kat.materialize(c, kt, vt)
    G, kout, aout = orc.groupby([keys], [vals], [(0, c["agg"], c["ddof"])], include_null_keys=c["include_null_keys"])
    kat.check(name, G, kout[0], aout[0], exp)


ARG_CASES = kat.load_cases("reference_kat_arg.json")


@pytest.mark.parametrize("name,c,kt,vt", ARG_CASES, ids=[x[0] for x in ARG_CASES])
def test_oracle_matches_reference_argmin_argmax_kat(name, c, kt, vt):
    """argmin_tests.cpp / argmax_tests.cpp: the result is the row number of the group's extremum (int32), null without a valid value"""
    keys, vals, exp = kat.materialize(c, kt, vt)
    G, kout, aout = orc.groupby([keys], [vals], [(0, c["agg"], c["ddof"])], include_null_keys=c["include_null_keys"])
    kat.check(name, G, kout[0], aout[0], exp)


@pytest.mark.parametrize("threads", [1, 4])
def test_oracle_threads_agree_with_numpy(threads):
    rng = np.random.default_rng(7)
    n = 300_000
    k = rng.integers(0, 1000, n).astype(np.int64)
    v = rng.integers(-1000, 1000, n).astype(np.int64)
    G, kout, aout = orc.groupby([orc.HostColumn(k)], [orc.HostColumn(v)], [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1)],
                                threads=threads)
    uk = np.unique(k)
    assert G == len(uk)
    np.testing.assert_array_equal(kout[0][0], uk)
    sums = np.zeros(1000, dtype=np.int64); np.add.at(sums, k, v)
    np.testing.assert_array_equal(aout[0][0], sums[uk])
    np.testing.assert_array_equal(aout[1][0], np.bincount(k, minlength=1000)[uk])
    mins = np.full(1000, np.iinfo(np.int64).max); np.minimum.at(mins, k, v)
    np.testing.assert_array_equal(aout[2][0], mins[uk])


def test_oracle_matches_pandas_two_keys_mean():
    pd = pytest.importorskip("pandas")
    rng = np.random.default_rng(42)
    n = 200_000
    a = rng.integers(0, 32, n).astype(np.int64); b = rng.integers(0, 32, n).astype(np.int64)
    c = rng.standard_normal(n); d = rng.standard_normal(n)
    G, kout, aout = orc.groupby([orc.HostColumn(a), orc.HostColumn(b)], [orc.HostColumn(c), orc.HostColumn(d)],
                                [(0, "mean", 1), (1, "mean", 1), (0, "variance", 1), (1, "count_valid", 1)])
    ref = pd.DataFrame({"A": a, "B": b, "C": c, "D": d}).groupby(["A", "B"]).agg(
        C=("C", "mean"), D=("D", "mean"), V=("C", "var"), N=("D", "count")).reset_index()
    assert G == len(ref)
    np.testing.assert_array_equal(kout[0][0], ref["A"].to_numpy())
    np.testing.assert_array_equal(kout[1][0], ref["B"].to_numpy())
    np.testing.assert_allclose(aout[0][0], ref["C"].to_numpy(), rtol=1e-9)
    np.testing.assert_allclose(aout[1][0], ref["D"].to_numpy(), rtol=1e-9)
    np.testing.assert_allclose(aout[2][0], ref["V"].to_numpy(), rtol=1e-9)
    np.testing.assert_array_equal(aout[3][0], ref["N"].to_numpy())


def test_oracle_rejects_invalid():
    k = orc.HostColumn(np.array([1, 2, 3], dtype=np.int32))
    with pytest.raises(orc.OracleError):  # groupby.cu:225-229 size mismatch
        orc.groupby([k], [orc.HostColumn(np.arange(5, dtype=np.int32))], [(0, "count_valid", 1)])
    ts = orc.HostColumn(np.arange(3, dtype=np.int64), type_id=13)
    with pytest.raises(orc.OracleError):  # timestamps cannot be summed (aggregation.hpp target_type void)
        orc.groupby([k], [ts], [(0, "sum", 1)])
