"""GPU tests of the multi-GPU path: owner-rank partition, validity repacking, merge kinds, moment
finalisation, and DistributedGroupBy end to end.

On a 1-GPU box the ranks are simulated as THREADS of one process sharing the GPU through a thread communicator, so
every kernel and every line of DistributedGroupBy.aggregate runs; the NCCL communicator itself is covered by the
2-process test at the bottom (skipped when fewer than 2 GPUs are visible) and by the gloo tests in test_distributed_gloo.py.
"""
import ctypes as C
import threading

import numpy as np
import pytest
import torch

import cudf_b200 as cb
from cudf_b200 import _lib
from cudf_b200 import aggregation as agg
from cudf_b200 import distributed as D
from oracle import oracle as orc
from tests.gpu_util import make_agg, sort_order, to_device

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------------------------------------------------------------
# building blocks


@pytest.mark.parametrize("n", [0, 1, 31, 32, 33, 1000, 100_003])
def test_bytes_to_bitmask(n):
    rng = np.random.default_rng(n)
    valid = rng.random(n) > 0.3
    col = D.column_from_wire(torch.zeros(n, dtype=torch.int64, device="cuda"), torch.from_numpy(valid.astype(np.uint8)).cuda(), 4)
    _, v = col.to_numpy()
    np.testing.assert_array_equal(v, valid)
    assert col.null_count == int(n - np.count_nonzero(valid))


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("n", [0, 1, 777, 200_000])
def test_partition_by_owner(world, n):
    rng = np.random.default_rng(7 * world + n)
    keys = rng.integers(-50, 50, n).astype(np.int64)
    vals = rng.standard_normal(n)
    valid = rng.random(n) > 0.25
    small = rng.integers(0, 100, n).astype(np.int16)
    kc = cb.Column.from_numpy(keys)
    cols = [kc, cb.Column.from_numpy(vals, valid), cb.Column.from_numpy(small)]
    bufs, off = D.partition_by_owner(cb.Table([kc]), cols, world)
    assert off[0] == 0 and off[world] == n and all(off[i] <= off[i + 1] for i in range(world))
    pk = bufs[0][0].cpu().numpy()
    pv = bufs[1][0].cpu().numpy()
    pvalid = bufs[1][1].cpu().numpy().astype(bool)
    ps = bufs[2][0].cpu().numpy()
    assert bufs[0][1] is None and bufs[2][1] is None
    # a key lives in exactly one run
    owner_of = {}
    for r in range(world):
        for k in np.unique(pk[off[r]:off[r + 1]]):
            assert owner_of.setdefault(int(k), r) == r
    # rows are permuted as whole rows (multiset equality on (key, value bits, valid, small))
    def rows(k, v, m, s):
        vb = np.where(m, v, 0.0).view(np.int64)
        a = np.stack([k, vb, m.astype(np.int64), s.astype(np.int64)], axis=1)
        return a[np.lexsort(a.T[::-1])]
    np.testing.assert_array_equal(rows(pk, pv, pvalid, ps), rows(keys, vals, valid, small))
    if world > 1 and n >= 200_000:
        sizes = np.diff(off)
        assert sizes.min() > 0  # 100 distinct keys spread over every owner


def test_partition_null_keys_and_two_key_columns():
    rng = np.random.default_rng(3)
    n = 5000
    k0 = rng.integers(0, 20, n).astype(np.int32)
    k0v = rng.random(n) > 0.1
    k1 = rng.integers(0, 3, n).astype(np.int8)
    c0, c1 = cb.Column.from_numpy(k0, k0v), cb.Column.from_numpy(k1)
    bufs, off = D.partition_by_owner(cb.Table([c0, c1]), [c0, c1], 4)
    a = bufs[0][0].cpu().numpy()
    av = bufs[0][1].cpu().numpy().astype(bool)
    b = bufs[1][0].cpu().numpy()
    seen = {}
    for r in range(4):
        for i in range(off[r], off[r + 1]):
            key = (int(a[i]) if av[i] else None, int(b[i]))
            assert seen.setdefault(key, r) == r
    assert int(av.sum()) == int(k0v.sum())


def test_merge_count_kind():
    keys = np.array([1, 2, 1, 3, 2, 1], dtype=np.int64)
    counts = np.array([5, 7, 11, 0, 1, 2], dtype=np.int32)
    gb = cb.GroupBy(cb.Table([cb.Column.from_numpy(keys)]))
    kt, rt = gb.aggregate([cb.GroupByRequest(cb.Column.from_numpy(counts), [agg.merge_count(), agg.sum()])])
    k, _ = kt.columns()[0].to_numpy()
    c, cv = rt[0].columns()[0].to_numpy()
    s, _ = rt[0].columns()[1].to_numpy()
    order = np.argsort(k)
    assert c.dtype == np.int32 and rt[0].columns()[0].mask is None
    np.testing.assert_array_equal(k[order], [1, 2, 3])
    np.testing.assert_array_equal(c[order], [18, 8, 0])
    np.testing.assert_array_equal(s[order], [18, 8, 0])
    with pytest.raises(cb.LogicError):
        gb.aggregate([cb.GroupByRequest(cb.Column.from_numpy(keys), [agg.merge_count()])])  # only INT32 counts merge


@pytest.mark.parametrize("dtype", [np.int64, np.float64, np.float32])
def test_finalize_moments_matches_single_gpu_finaliser(dtype):
    rng = np.random.default_rng(11)
    n = 4000
    keys = rng.integers(0, 37, n).astype(np.int64)
    vals = rng.integers(-100, 100, n).astype(dtype)
    valid = rng.random(n) > 0.5
    valid[keys == 5] = False  # one group without a valid value
    kc, vc = cb.Column.from_numpy(keys), cb.Column.from_numpy(vals, valid)
    gb = cb.GroupBy(cb.Table([kc]))
    kt, rt = gb.aggregate([cb.GroupByRequest(vc, [agg.sum(), agg.sum_of_squares(), agg.count(), agg.mean(), agg.m2(), agg.variance(1),
                                                  agg.std(0)])])
    s, q, c, mean, m2, var, std = rt[0].columns()
    for kind, ddof, ref, with_mask in ((agg.Kind.MEAN, 0, mean, True), (agg.Kind.M2, 0, m2, False), (agg.Kind.VARIANCE, 1, var, True),
                                       (agg.Kind.STD, 0, std, True)):
        out = D.finalize_moments(kind, ddof, s, None if kind == agg.Kind.MEAN else q, c, with_mask)
        d, v = out.to_numpy()
        rd, rv = ref.to_numpy()
        np.testing.assert_array_equal(v, rv)
        np.testing.assert_array_equal(d[v].view(np.int64), rd[rv].view(np.int64))  # same arithmetic -> same bits
        assert out.null_count == int((~v).sum())


# ---------------------------------------------------------------------------------------------------------------------
# DistributedGroupBy end to end, ranks simulated as threads sharing cuda:0


class _Shared:
    def __init__(self, world):
        self.world = world
        self.barrier = threading.Barrier(world)
        self.vecs = [None] * world
        self.bufs = [None] * world
        self.ctxs = [None] * world


class ThreadComm:
    """Same interface as distributed.TorchComm; the exchange is device-side slicing + concatenation between threads."""

    def __init__(self, shared, rank):
        self.s, self.rank, self.world = shared, rank, shared.world

    def gather_counts(self, vec, device):
        self.s.vecs[self.rank] = list(vec)
        self.s.barrier.wait()
        m = torch.tensor(self.s.vecs, dtype=torch.int64)
        self.s.barrier.wait()
        return m

    def peer_connect(self, region_bytes, device):
        """Same-process ranks: map each other's exchange buffers directly (bgb_peer_attach_local)."""
        ctx = D.PeerContext(self.rank, self.world, region_bytes, device)
        self.s.ctxs[self.rank] = ctx
        self.s.barrier.wait()
        for r in range(self.world):
            if r != self.rank:
                ctx.attach_local(r, self.s.ctxs[r])
        self.s.barrier.wait()
        return ctx

    def all_to_all(self, buffers, send_counts, recv_counts):
        torch.cuda.synchronize()
        off = np.concatenate([[0], np.cumsum(send_counts)])
        self.s.bufs[self.rank] = (buffers, off)
        self.s.barrier.wait()
        out = []
        for i, b in enumerate(buffers):
            if b is None:
                assert all(self.s.bufs[r][0][i] is None for r in range(self.world)), "validity flag disagrees across ranks"
                out.append(None)
                continue
            parts = []
            for src in range(self.world):
                sb, so = self.s.bufs[src]
                parts.append(sb[i][int(so[self.rank]):int(so[self.rank + 1])])
                assert parts[-1].numel() == recv_counts[src]
            out.append(torch.cat(parts))
        torch.cuda.synchronize()
        self.s.barrier.wait()
        return out


EXCHANGES = ["nccl", "peer"]
LAST = {}


def run_simulated(world, shards, aggs, include_null_keys=False, options=None, exchange="nccl", peer_region_bytes=8 << 20):
    """shards: per rank (keys [HostColumn], values [HostColumn]). aggs: [(value_index, kind, ddof)].
    Returns key-sorted (G, keys, results) over all ranks plus per-rank key sets.
    Every simulated rank runs on its own CUDA stream: the peer path's wait kernel spins until the OTHER ranks' scatter
    kernels have run, which on one GPU needs them to be able to run concurrently."""
    shared = _Shared(world)
    results, errors = [None] * world, []
    comms = [ThreadComm(shared, r) for r in range(world)]

    def work(rank):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                work_on_stream(rank)
        except BaseException as e:  # noqa: BLE001
            errors.append(e)
            shared.barrier.abort()

    def work_on_stream(rank):
        if True:
            keys, values = shards[rank]
            dk = [to_device(k) for k in keys]
            dv = [to_device(v) for v in values]
            per_val = {}
            for pos, (vi, kind, ddof) in enumerate(aggs):
                per_val.setdefault(vi, []).append((pos, kind, ddof))
            reqs, where = [], {}
            for vi, lst in per_val.items():
                reqs.append(cb.GroupByRequest(dv[vi], [make_agg(k, d) for _, k, d in lst]))
                for j, (pos, _, _) in enumerate(lst):
                    where[pos] = (len(reqs) - 1, j)
            dgb = D.DistributedGroupBy(cb.Table(dk), cb.NullPolicy.INCLUDE if include_null_keys else cb.NullPolicy.EXCLUDE,
                                       options=options, comm=comms[rank], exchange=exchange, peer_region_bytes=peer_region_bytes)
            kt, rts = dgb.aggregate(reqs)
            torch.cuda.current_stream().synchronize()
            kout = [c.to_numpy() for c in kt.columns()]
            aout = []
            for pos in range(len(aggs)):
                r, j = where[pos]
                aout.append(rts[r].columns()[j].to_numpy())
            results[rank] = (kout, aout)
            LAST[rank] = (dgb.last_exchange, dgb.last_path)

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for c in shared.ctxs:
        if c is not None:
            c.close()
    for c in comms:
        ctx = getattr(c, "_bgb_peer_ctx", None)
        if ctx is not None and ctx not in shared.ctxs:
            ctx.close()
    if errors:
        raise errors[0]
    nk = len(results[0][0])
    kall = [(np.concatenate([results[r][0][c][0] for r in range(world)]), np.concatenate([results[r][0][c][1] for r in range(world)]))
            for c in range(nk)]
    aall = [(np.concatenate([results[r][1][a][0] for r in range(world)]), np.concatenate([results[r][1][a][1] for r in range(world)]))
            for a in range(len(aggs))]
    order = sort_order(kall)
    per_rank_sizes = [len(results[r][0][0][0]) for r in range(world)]
    return len(order), [(d[order], v[order]) for d, v in kall], [(d[order], v[order]) for d, v in aall], per_rank_sizes


def concat_shards(shards):
    def cat(cols):
        data = np.concatenate([c.data for c in cols])
        if any(c.valid is not None for c in cols):
            valid = np.concatenate([c.valid if c.valid is not None else np.ones(len(c.data), bool) for c in cols])
        else:
            valid = None
        return orc.HostColumn(data, valid, cols[0].type_id)
    nk, nv = len(shards[0][0]), len(shards[0][1])
    return [cat([s[0][i] for s in shards]) for i in range(nk)], [cat([s[1][i] for s in shards]) for i in range(nv)]


def check_against_oracle(world, shards, aggs, include_null_keys=False, options=None, rtol=1e-9, exchange="nccl", **kw):
    keys, values = concat_shards(shards)
    Go, ko, ao = orc.groupby(keys, values, aggs, include_null_keys=include_null_keys, threads=4)
    G, kg, ag, sizes = run_simulated(world, shards, aggs, include_null_keys, options, exchange=exchange, **kw)
    assert all(LAST[r][0] == exchange for r in range(world)), LAST
    assert G == Go, f"{G} groups over {world} ranks, oracle has {Go}"
    for c in range(len(keys)):
        np.testing.assert_array_equal(kg[c][1], ko[c][1])
        np.testing.assert_array_equal(kg[c][0][kg[c][1]], ko[c][0][ko[c][1]])
    for a in range(len(aggs)):
        np.testing.assert_array_equal(ag[a][1], ao[a][1], err_msg=f"agg {aggs[a]} validity")
        got, exp = ag[a][0][ag[a][1]], ao[a][0][ao[a][1]]
        if got.dtype.kind == "f":
            np.testing.assert_allclose(got, exp, rtol=rtol, atol=0, err_msg=f"agg {aggs[a]}")  # north_star: 1e-9 relative
        else:
            np.testing.assert_array_equal(got, exp, err_msg=f"agg {aggs[a]}")
    return sizes


def int_shards(world, n, groups, null_frac=0.0, seed=0, key_dtype=np.int64):
    out = []
    for r in range(world):
        rng = np.random.default_rng(seed + 17 * r)
        k = rng.integers(0, groups, n).astype(key_dtype)
        v = rng.integers(-10_000, 10_000, n).astype(np.int64)
        valid = (rng.random(n) >= null_frac) if null_frac > 0 else None
        out.append(([orc.HostColumn(k)], [orc.HostColumn(v, valid)]))
    return out


@pytest.mark.parametrize("world", [1, 2, 3, 4])
@pytest.mark.parametrize("exchange", EXCHANGES)
def test_sum_count_int64(exchange, world):
    shards = int_shards(world, 100_000, 1000)
    sizes = check_against_oracle(world, shards, [(0, "sum", 0), (0, "count_valid", 0), (0, "count_all", 0)], exchange=exchange)
    assert sum(sizes) == 1000 and (world == 1 or min(sizes) > 0)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("exchange", EXCHANGES)
def test_all_kinds_with_nulls(exchange, world):
    shards = int_shards(world, 50_000, 500, null_frac=0.3, seed=5)
    check_against_oracle(world, shards, [(0, "sum", 0), (0, "min", 0), (0, "max", 0), (0, "count_valid", 0), (0, "mean", 0),
                                         (0, "variance", 1), (0, "std", 0), (0, "m2", 0), (0, "sum_of_squares", 0), (0, "count_all", 0)], exchange=exchange)


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_only_one_rank_has_a_null_mask(exchange):
    shards = int_shards(2, 20_000, 300, null_frac=0.0, seed=9)
    k, v = shards[0]
    valid = np.random.default_rng(1).random(20_000) > 0.5
    shards[0] = (k, [orc.HostColumn(v[0].data, valid)])
    check_against_oracle(2, shards, [(0, "sum", 0), (0, "max", 0), (0, "mean", 0)], exchange=exchange)


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_group_null_on_every_rank_stays_null(exchange):
    shards = int_shards(2, 5000, 50, null_frac=0.2, seed=2)
    for k, v in shards:
        v[0].valid[k[0].data == 7] = False
    check_against_oracle(2, shards, [(0, "sum", 0), (0, "min", 0), (0, "mean", 0), (0, "variance", 1), (0, "count_valid", 0)], exchange=exchange)


@pytest.mark.parametrize("include", [False, True])
@pytest.mark.parametrize("exchange", EXCHANGES)
def test_null_keys_two_key_columns(exchange, include):
    shards = []
    for r in range(3):
        rng = np.random.default_rng(40 + r)
        n = 30_000
        k0 = rng.integers(0, 40, n).astype(np.int32)
        k0v = rng.random(n) > 0.1
        k1 = rng.integers(0, 4, n).astype(np.int8)
        v = rng.standard_normal(n) * 100
        vv = rng.random(n) > 0.1
        shards.append(([orc.HostColumn(k0, k0v), orc.HostColumn(k1)], [orc.HostColumn(v, vv)]))
    check_against_oracle(3, shards, [(0, "sum", 0), (0, "min", 0), (0, "max", 0), (0, "mean", 0), (0, "std", 1), (0, "count_valid", 0)],
                         include_null_keys=include, exchange=exchange)


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_float_values_two_requests(exchange):
    shards = []
    for r in range(2):
        rng = np.random.default_rng(60 + r)
        n = 80_000
        k = rng.integers(0, 5000, n).astype(np.int64)
        a = rng.standard_normal(n)
        b = rng.integers(0, 1000, n).astype(np.int32)
        shards.append(([orc.HostColumn(k)], [orc.HostColumn(a), orc.HostColumn(b)]))
    check_against_oracle(2, shards, [(0, "sum", 0), (1, "sum", 0), (0, "mean", 0), (1, "max", 0), (1, "variance", 1), (0, "min", 0)], exchange=exchange)


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_keys_only_and_empty_rank(exchange):
    shards = int_shards(2, 10_000, 100, seed=3)
    # rank 1 holds no rows at all
    shards[1] = ([orc.HostColumn(np.empty(0, np.int64))], [orc.HostColumn(np.empty(0, np.int64))])
    check_against_oracle(2, shards, [(0, "sum", 0), (0, "count_valid", 0)], exchange=exchange)
    check_against_oracle(2, int_shards(2, 10_000, 100, seed=3), [], exchange=exchange)


@pytest.mark.parametrize("exchange", EXCHANGES)
def test_high_cardinality_goes_through_the_partitioned_merge(exchange):
    shards = int_shards(2, 400_000, 300_000, seed=8)
    sizes = check_against_oracle(2, shards, [(0, "sum", 0), (0, "count_valid", 0)], exchange=exchange)
    assert abs(sizes[0] - sizes[1]) < 0.05 * sum(sizes)


@pytest.mark.parametrize("world", [2, 3])
@pytest.mark.parametrize("aggs", [[(0, "sum", 0), (0, "count_valid", 0)],
                                  [(0, "min", 0), (0, "max", 0)],
                                  [(0, "mean", 0), (0, "variance", 1)],
                                  [(0, "sum", 0), (0, "min", 0), (0, "max", 0), (0, "count_all", 0)]], ids=["lean", "minmax", "moments", "five-cells"])
def test_peer_fused_exchange_and_slice_merge(monkeypatch, world, aggs):
    """The records-form local build writes every finished slice straight into the owners' regions (no separate scatter pass over
    the local groups) and the owner merges what it received through the same slice kernels (no HBM hash table, no compaction)."""
    monkeypatch.setenv("BGB_HC_MERGE_MIN", "1000")
    nulls = 0.0 if aggs[0][1] == "sum" and len(aggs) == 2 else 0.2
    shards = int_shards(world, 1_500_000, 700_000, null_frac=nulls, seed=21)
    # one all-ones key per rank: its group lives in the reserved slot of the slices and travels through hc_finish
    for k, _ in shards:
        k[0].data[::100_003] = -1
    sizes = check_against_oracle(world, shards, aggs, options={"force_path": 7}, exchange="peer", peer_region_bytes=64 << 20)
    # records of more than three cells do not fit the 32-byte record of the slice kernels: merged in the HBM table
    want = "bucketed+peer-fused" if len(aggs) == 4 else "bucketed+peer-fused+slices"
    assert all(LAST[r][1] == want for r in range(world)), LAST
    assert abs(sizes[0] - sizes[1]) < 0.05 * sum(sizes)


def test_peer_slice_merge_after_the_separate_scatter(monkeypatch):
    """BGB_NO_PEER_FUSE: the local groups are scattered by their own kernel; the merge still runs through the slices."""
    monkeypatch.setenv("BGB_HC_MERGE_MIN", "1000")
    monkeypatch.setenv("BGB_NO_PEER_FUSE", "1")
    shards = int_shards(2, 300_000, 200_000, seed=23)
    check_against_oracle(2, shards, [(0, "sum", 0), (0, "count_valid", 0)], exchange="peer", peer_region_bytes=64 << 20)
    assert all(LAST[r][1].endswith("+peer+slices") for r in range(2)), LAST


def test_peer_region_grows_after_overflow():
    """A 4 KB region holds 64 records per (source, destination) segment; 1000 groups overflow it on every rank, every rank
    reports BGB_PEER_OVERFLOW in the same step and the communicator re-creates a larger region until the exchange fits."""
    shards = int_shards(2, 50_000, 1000, seed=12)
    check_against_oracle(2, shards, [(0, "sum", 0), (0, "count_valid", 0)], exchange="peer", peer_region_bytes=4096)


def test_peer_two_steps_reuse_the_context():
    """Consecutive calls alternate between the two regions (parity of the epoch) without any barrier in between."""
    world = 2
    shared = _Shared(world)
    comms = [ThreadComm(shared, r) for r in range(world)]
    out, errors = {}, []

    def work(rank):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                for step in range(5):
                    rng = np.random.default_rng(1000 * step + rank)
                    k = rng.integers(0, 200 + 50 * step, 20_000).astype(np.int64)
                    v = rng.integers(-5, 5, 20_000).astype(np.int64)
                    dgb = D.DistributedGroupBy(cb.Table([cb.Column.from_numpy(k)]), comm=comms[rank], exchange="peer", peer_region_bytes=1 << 20)
                    kt, rts = dgb.aggregate([cb.GroupByRequest(cb.Column.from_numpy(v), [agg.sum(), agg.count()])])
                    torch.cuda.current_stream().synchronize()
                    out[(step, rank)] = (k, v, kt.columns()[0].to_numpy()[0], rts[0].columns()[0].to_numpy()[0], rts[0].columns()[1].to_numpy()[0])
        except BaseException as e:  # noqa: BLE001
            errors.append(e)
            shared.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for c in shared.ctxs:
        if c is not None:
            c.close()
    if errors:
        raise errors[0]
    for step in range(5):
        k = np.concatenate([out[(step, r)][0] for r in range(world)])
        v = np.concatenate([out[(step, r)][1] for r in range(world)])
        gk = np.concatenate([out[(step, r)][2] for r in range(world)])
        gs = np.concatenate([out[(step, r)][3] for r in range(world)])
        gc = np.concatenate([out[(step, r)][4] for r in range(world)])
        order = np.argsort(gk)
        uk, inv = np.unique(k, return_inverse=True)
        np.testing.assert_array_equal(gk[order], uk)
        np.testing.assert_array_equal(gs[order], np.bincount(inv, weights=v).astype(np.int64))
        np.testing.assert_array_equal(gc[order], np.bincount(inv))


def test_peer_merge_size_follows_the_last_step_and_survives_a_wrong_guess(monkeypatch):
    """The slices of the merge are sized from the previous step's groups-per-record ratio (groupby.cu: merge_ratio). Steps 0-1: every
    rank sees every key (ratio 1/world, the second step runs on small slices); step 2: the ranks' key sets are DISJOINT (one group
    per record: the small slices fill, the overflow flag reruns the merge at the safe size); step 3: shared keys again."""
    monkeypatch.setenv("BGB_HC_MERGE_MIN", "1000")
    # sized so that step 2 MUST overflow: steps 0-1 leave a ratio of ~0.27 groups per record, which sizes step 2's slices for ~1.05e6
    # slots while its ~1.47e6 records per rank are all distinct
    world, n, K = 4, 4_000_000, 1_600_000
    shared = _Shared(world)
    comms = [ThreadComm(shared, r) for r in range(world)]
    out, errors = {}, []

    def keys_of(step, rank):
        rng = np.random.default_rng(77 * step + rank)
        if step == 2:
            return rng.integers(0, K, n).astype(np.int64) * world + rank  # disjoint over the ranks
        return rng.integers(0, K, n).astype(np.int64)

    def work(rank):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                for step in range(4):
                    k = keys_of(step, rank)
                    v = np.random.default_rng(5 + step + 10 * rank).integers(-5, 5, n).astype(np.int64)
                    dgb = D.DistributedGroupBy(cb.Table([cb.Column.from_numpy(k)]), comm=comms[rank], exchange="peer", peer_region_bytes=64 << 20,
                                               options={"force_path": 7})
                    kt, rts = dgb.aggregate([cb.GroupByRequest(cb.Column.from_numpy(v), [agg.sum(), agg.count()])])
                    torch.cuda.current_stream().synchronize()
                    assert dgb.last_path == "bucketed+peer-fused+slices", dgb.last_path
                    out[(step, rank)] = (k, v, kt.columns()[0].to_numpy()[0], rts[0].columns()[0].to_numpy()[0], rts[0].columns()[1].to_numpy()[0])
        except BaseException as e:  # noqa: BLE001
            errors.append(e)
            shared.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for c in shared.ctxs:
        if c is not None:
            c.close()
    if errors:
        raise errors[0]
    for step in range(4):
        k = np.concatenate([out[(step, r)][0] for r in range(world)])
        v = np.concatenate([out[(step, r)][1] for r in range(world)])
        gk = np.concatenate([out[(step, r)][2] for r in range(world)])
        gs = np.concatenate([out[(step, r)][3] for r in range(world)])
        gc = np.concatenate([out[(step, r)][4] for r in range(world)])
        order = np.argsort(gk)
        uk, inv = np.unique(k, return_inverse=True)
        np.testing.assert_array_equal(gk[order], uk)
        np.testing.assert_array_equal(gs[order], np.bincount(inv, weights=v).astype(np.int64))
        np.testing.assert_array_equal(gc[order], np.bincount(inv))


def test_peer_rank_failure_is_broadcast_and_the_context_survives():
    """One rank fails before it can scatter (its value column is shorter than its keys: cudf::logic_error, groupby.cu:186-229): it
    still signals -- with the abort word -- so the other rank returns BGB_PEER_ABORTED in the SAME step instead of spinning until
    the 20 s time-out, both ranks are at the same epoch afterwards, and the next step on the same contexts is correct
    (round-1 advisor finding: a single failing rank desynchronised the exchange for good)."""
    import time
    world = 2
    shared = _Shared(world)
    comms = [ThreadComm(shared, r) for r in range(world)]
    out, raised, errors = {}, {}, []

    def work(rank):
        try:
            with torch.cuda.stream(torch.cuda.Stream()):
                for step in range(3):
                    rng = np.random.default_rng(77 * step + rank)
                    k = rng.integers(0, 300, 10_000).astype(np.int64)
                    v = rng.integers(-5, 5, 10_000).astype(np.int64)
                    vcol = cb.Column.from_numpy(v[:-1] if (step == 1 and rank == 1) else v)   # step 1, rank 1: size mismatch
                    dgb = D.DistributedGroupBy(cb.Table([cb.Column.from_numpy(k)]), comm=comms[rank], exchange="peer", peer_region_bytes=1 << 20)
                    t0 = time.perf_counter()
                    try:
                        kt, rts = dgb.aggregate([cb.GroupByRequest(vcol, [agg.sum(), agg.count()])])
                        torch.cuda.current_stream().synchronize()
                        out[(step, rank)] = (k, v, kt.columns()[0].to_numpy()[0], rts[0].columns()[0].to_numpy()[0])
                    except _lib.BgbError as e:
                        raised[(step, rank)] = (e, time.perf_counter() - t0)
        except BaseException as e:  # noqa: BLE001
            errors.append(e)
            shared.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for c in shared.ctxs:
        if c is not None:
            c.close()
    if errors:
        raise errors[0]
    assert set(raised) == {(1, 0), (1, 1)}, raised
    assert isinstance(raised[(1, 1)][0], cb.LogicError)                       # the failing rank reports its own error
    assert isinstance(raised[(1, 0)][0], _lib.PeerAborted), raised[(1, 0)][0]   # its peer is told, in the same step
    assert raised[(1, 0)][1] < 5.0, "the healthy rank waited for the time-out"
    for step in (0, 2):                                                         # the steps around the failure are exact
        k = np.concatenate([out[(step, r)][0] for r in range(world)])
        v = np.concatenate([out[(step, r)][1] for r in range(world)])
        gk = np.concatenate([out[(step, r)][2] for r in range(world)])
        gs = np.concatenate([out[(step, r)][3] for r in range(world)])
        order = np.argsort(gk)
        uk, inv = np.unique(k, return_inverse=True)
        np.testing.assert_array_equal(gk[order], uk)
        np.testing.assert_array_equal(gs[order], np.bincount(inv, weights=v).astype(np.int64))


@pytest.mark.parametrize("exchange", EXCHANGES)
@pytest.mark.parametrize("nshards,null_frac", [(2, 0.0), (3, 0.3), (5, 0.0)])
def test_aggregate_shards_on_one_rank(exchange, nshards, null_frac):
    # several shards per rank (strong scaling of BASELINE configs[4]): per-shard partials folded on the rank, then the ordinary
    # distributed merge; the result is the groupby of the concatenated shards (the reference's verify_against_groupby idea,
    # cpp/tests/groupby/streaming_groupby_test.cpp:60-110)
    shards = int_shards(nshards, 20_000, 700, null_frac=null_frac, seed=5)
    if null_frac:
        shards[1][1][0].valid[shards[1][0][0].data == 3] = False  # a group that is all-null in one shard only
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (0, "max", 1), (0, "mean", 1), (0, "variance", 1),
            (0, "std", 0), (0, "m2", 1)]
    keys, values = concat_shards(shards)
    Go, ko, ao = orc.groupby(keys, values, aggs, threads=4)
    dev_shards = []
    for k, v in shards:
        dk, dv = [to_device(c) for c in k], [to_device(c) for c in v]
        dev_shards.append((cb.Table(dk), [cb.GroupByRequest(dv[0], [make_agg(kind, d) for _, kind, d in aggs])]))
    dgb = D.DistributedGroupBy(dev_shards[0][0], comm=D.SingleComm(), exchange=exchange, peer_region_bytes=8 << 20)
    kt, rts = dgb.aggregate_shards(dev_shards)
    ctx = getattr(dgb._comm, "_bgb_peer_ctx", None)
    kd, kv = kt.columns()[0].to_numpy()
    order = np.argsort(kd)
    assert len(kd) == Go
    np.testing.assert_array_equal(kd[order], ko[0][0])
    for a in range(len(aggs)):
        d, v = rts[0].columns()[a].to_numpy()
        np.testing.assert_array_equal(v[order], ao[a][1], err_msg=f"agg {aggs[a]} validity")
        got, exp = d[order][ao[a][1]], ao[a][0][ao[a][1]]
        if got.dtype.kind == "f":
            np.testing.assert_allclose(got, exp, rtol=1e-9, atol=0, err_msg=f"agg {aggs[a]}")
        else:
            np.testing.assert_array_equal(got, exp, err_msg=f"agg {aggs[a]}")
    if ctx is not None:
        ctx.close()


def test_unsupported_kind_raises():
    k = cb.Column.from_numpy(np.arange(10, dtype=np.int64))
    dgb = D.DistributedGroupBy(cb.Table([k]), comm=D.SingleComm())
    with pytest.raises(cb.UnsupportedError):
        dgb.aggregate([cb.GroupByRequest(k, [agg.argmax()])])


# ---------------------------------------------------------------------------------------------------------------------
# the real thing: 2 processes, 2 GPUs, NCCL


def _nccl_worker(rank, world, port, outdir):
    import os
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        keys, values = int_shards(world, 200_000, 20_000, null_frac=0.1, seed=77)[rank]
        dk = [cb.Column.from_numpy(k.data, k.valid, device=f"cuda:{rank}") for k in keys]
        dv = [cb.Column.from_numpy(v.data, v.valid, device=f"cuda:{rank}") for v in values]
        for exchange in ("nccl", "peer"):
            dgb = D.DistributedGroupBy(cb.Table(dk), exchange=exchange)
            kt, rts = dgb.aggregate([cb.GroupByRequest(dv[0], [agg.sum(), agg.count(), agg.min(), agg.mean()])])
            assert dgb.last_exchange == exchange
            cols = [kt.columns()[0].to_numpy()] + [c.to_numpy() for c in rts[0].columns()]
            np.savez(os.path.join(outdir, f"{exchange}_r{rank}.npz"), **{f"d{i}": c[0] for i, c in enumerate(cols)},
                     **{f"v{i}": c[1] for i, c in enumerate(cols)})
        # the fused exchange (slices write into the other PROCESS's region through the CUDA IPC mapping) + the slice merge
        os.environ["BGB_HC_MERGE_MIN"] = "1000"
        keys, values = int_shards(world, 1_500_000, 700_000, seed=21)[rank]
        dk = [cb.Column.from_numpy(k.data, k.valid, device=f"cuda:{rank}") for k in keys]
        dv = [cb.Column.from_numpy(v.data, v.valid, device=f"cuda:{rank}") for v in values]
        dgb = D.DistributedGroupBy(cb.Table(dk), exchange="peer", options={"force_path": 7}, peer_region_bytes=64 << 20)
        for step in range(2):  # twice: both parities of the double-buffered regions
            kt, rts = dgb.aggregate([cb.GroupByRequest(dv[0], [agg.sum(), agg.count()])])
        assert dgb.last_path == "bucketed+peer-fused+slices", dgb.last_path
        cols = [kt.columns()[0].to_numpy()] + [c.to_numpy() for c in rts[0].columns()]
        np.savez(os.path.join(outdir, f"fused_r{rank}.npz"), **{f"d{i}": c[0] for i, c in enumerate(cols)}, **{f"v{i}": c[1] for i, c in enumerate(cols)})
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_nccl_two_gpus(tmp_path):
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    world = 2
    mp.spawn(_nccl_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    shards = int_shards(world, 200_000, 20_000, null_frac=0.1, seed=77)
    keys, values = concat_shards(shards)
    Go, ko, ao = orc.groupby(keys, values, [(0, "sum", 0), (0, "count_valid", 0), (0, "min", 0), (0, "mean", 0)], threads=4)
    for exchange in ("nccl", "peer"):
        parts = [np.load(tmp_path / f"{exchange}_r{r}.npz") for r in range(world)]
        k = np.concatenate([p["d0"] for p in parts])
        order = np.argsort(k, kind="stable")
        assert len(k) == Go
        np.testing.assert_array_equal(k[order], ko[0][0])
        for i in range(4):
            d = np.concatenate([p[f"d{i + 1}"] for p in parts])[order]
            v = np.concatenate([p[f"v{i + 1}"] for p in parts])[order]
            np.testing.assert_array_equal(v, ao[i][1])
            if d.dtype.kind == "f":
                np.testing.assert_allclose(d[v], ao[i][0][ao[i][1]], rtol=1e-9, atol=0)
            else:
                np.testing.assert_array_equal(d[v], ao[i][0][ao[i][1]])
    keys, values = concat_shards(int_shards(world, 1_500_000, 700_000, seed=21))
    Go, ko, ao = orc.groupby(keys, values, [(0, "sum", 0), (0, "count_valid", 0)], threads=4)
    parts = [np.load(tmp_path / f"fused_r{r}.npz") for r in range(world)]
    k = np.concatenate([p["d0"] for p in parts])
    order = np.argsort(k, kind="stable")
    assert len(k) == Go
    np.testing.assert_array_equal(k[order], ko[0][0])
    for i in range(2):
        np.testing.assert_array_equal(np.concatenate([p[f"d{i + 1}"] for p in parts])[order], ao[i][0])
