"""NDS-H query 1 through the hash groupby (reference driver /root/reference/cpp/benchmarks/ndsh/q01.cpp:
101-152): the query's ten result columns against a plain numpy evaluation of the SQL on the same synthetic lineitem."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from cudf_b200 import ndsh  # noqa: E402


def _numpy_q1(c, cutoff):
    keep = c["l_shipdate"] <= cutoff
    rf, ls = c["l_returnflag"][keep], c["l_linestatus"][keep]
    qty, price, disc, tax = (c[k][keep] for k in ("l_quantity", "l_extendedprice", "l_discount", "l_tax"))
    disc_price = price * (1.0 - disc)
    charge = disc_price * (1.0 + tax)
    code = rf.astype(np.int64) * 4 + ls
    groups = np.unique(code)
    out = {k: [] for k in ndsh.RESULT_COLUMNS}
    for g in groups:
        m = code == g
        out["l_returnflag"].append(g // 4); out["l_linestatus"].append(g % 4)
        out["sum_base_price"].append(price[m].sum()); out["avg_price"].append(price[m].sum() / m.sum())
        out["sum_qty"].append(qty[m].sum()); out["avg_qty"].append(qty[m].sum() / m.sum())
        out["avg_disc"].append(disc[m].sum() / m.sum())
        out["sum_disc_price"].append(disc_price[m].sum()); out["sum_charge"].append(charge[m].sum())
        out["count_order"].append(m.sum())
    return {k: np.asarray(v) for k, v in out.items()}


@pytest.mark.parametrize("n", [1000, 2_000_000])
def test_q1_matches_numpy(n):
    dev, host = ndsh.synthetic_lineitem(n, seed=7)
    got, path = ndsh.q1(dev)
    exp = _numpy_q1(host, ndsh.SHIPDATE_CUTOFF_DAYS)
    assert len(got["l_returnflag"]) == len(exp["l_returnflag"]) == 4
    for k in ("l_returnflag", "l_linestatus", "count_order"):
        np.testing.assert_array_equal(got[k].astype(np.int64), exp[k].astype(np.int64), err_msg=k)
    for k in ("sum_base_price", "sum_qty"):  # whole numbers < 2^53: exact in any order
        np.testing.assert_array_equal(got[k], exp[k], err_msg=k)
    for k in ("avg_price", "avg_qty", "avg_disc", "sum_disc_price", "sum_charge"):
        np.testing.assert_allclose(got[k], exp[k], rtol=1e-9, atol=0, err_msg=k)  # north_star: 1e-9 relative
    if n >= 65536:
        assert path.startswith("tiny"), path  # four groups: the tiny-cardinality kernel


def test_a_selective_cutoff():
    # only ~1000 of 300 000 rows pass: the filter is the key's null mask under EXCLUDE, so 99.7 % of the rows are dropped in-kernel
    dev, host = ndsh.synthetic_lineitem(300_000, seed=9)
    cutoff = int(np.sort(host["l_shipdate"])[1000])
    got, _ = ndsh.q1(dev, shipdate_cutoff_days=cutoff)
    exp = _numpy_q1(host, cutoff)
    assert len(got["count_order"]) == len(exp["count_order"])
    np.testing.assert_array_equal(got["count_order"].astype(np.int64), exp["count_order"].astype(np.int64))
    np.testing.assert_array_equal(got["sum_qty"], exp["sum_qty"])
