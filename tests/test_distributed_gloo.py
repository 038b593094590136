"""CPU tests (gloo, world_size 2) of the multi-GPU host logic in cudf_b200/distributed.py: decomposition of the
requested aggregations, the count-matrix all-gather and the all-to-all-v of partial columns.

The compute steps (local aggregation, owner partition, merge) are CUDA-only in the product; here the ORACLE stands in for
them as the checker, with `key % world` as a stand-in owner function -- the exchange logic under test does not depend on
which owner function permuted the partials. The GPU tests (test_gpu_distributed.py) run the real kernels end to end.
"""
import os
import socket

import numpy as np
import pytest
import torch

from cudf_b200 import aggregation as agg
from cudf_b200 import distributed as D
from cudf_b200._lib import UnsupportedError
from cudf_b200.aggregation import Kind


def test_decompose_simple_kinds():
    p = D.decompose([agg.sum(), agg.count(), agg.min(), agg.max()])
    assert p.partial_kinds == [Kind.SUM, Kind.COUNT_VALID, Kind.MIN, Kind.MAX]
    assert p.recipes == [("copy", 0), ("copy", 1), ("copy", 2), ("copy", 3)]


def test_decompose_mean_var_share_partials():
    # mean -> (sum, count); var/std -> (sum, sumsq, count): groupby.py:74-98,141-262 of cudf_polars' streaming decomposition
    p = D.decompose([agg.mean(), agg.variance(1), agg.std(0), agg.sum(), agg.m2()])
    assert p.partial_kinds == [Kind.SUM, Kind.COUNT_VALID, Kind.SUM_OF_SQUARES]
    assert p.recipes[0] == ("mean", 0, 1)
    assert p.recipes[1] == ("moment", Kind.VARIANCE, 1, 0, 2, 1)
    assert p.recipes[2] == ("moment", Kind.STD, 0, 0, 2, 1)
    assert p.recipes[3] == ("copy", 0)
    assert p.recipes[4][0:2] == ("moment", Kind.M2)


def test_decompose_duplicates_answered_twice():
    p = D.decompose([agg.sum(), agg.sum()])
    assert p.partial_kinds == [Kind.SUM] and p.recipes == [("copy", 0), ("copy", 0)]


def test_merge_kinds():
    assert D.merge_kind(Kind.SUM) == Kind.SUM
    assert D.merge_kind(Kind.SUM_OF_SQUARES) == Kind.SUM
    assert D.merge_kind(Kind.COUNT_VALID) == Kind.MERGE_COUNT
    assert D.merge_kind(Kind.COUNT_ALL) == Kind.MERGE_COUNT
    assert D.merge_kind(Kind.MIN) == Kind.MIN and D.merge_kind(Kind.MAX) == Kind.MAX


def test_decompose_rejects_row_index_aggregations():
    with pytest.raises(UnsupportedError):
        D.decompose([agg.argmin()])
    with pytest.raises(UnsupportedError):
        D.decompose([agg.median()])


def test_single_comm_is_identity():
    c = D.SingleComm()
    m = c.gather_counts([5, 1, 0], None)
    assert m.tolist() == [[5, 1, 0]]
    b = [torch.arange(5), None]
    assert c.all_to_all(b, [5], [5])[0] is b[0]


# ---------------------------------------------------------------------------------------------------------------------


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _shard(rank, n=5000, groups=300):
    rng = np.random.default_rng(100 + rank)
    keys = rng.integers(-groups // 2, groups // 2, n).astype(np.int64)
    vals = rng.integers(-1000, 1000, n).astype(np.int64)
    valid = rng.random(n) > 0.2
    if rank == 1:  # rank 1's column has no null mask at all: the validity flag must still agree across ranks
        valid = None
    return keys, vals, valid


def _worker(rank, world, port, outdir):
    import torch.distributed as dist
    from oracle import oracle as orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        keys, vals, valid = _shard(rank)
        # local partials (oracle as the stand-in for the CUDA local aggregation)
        G, ko, ao = orc.groupby([orc.HostColumn(keys)], [orc.HostColumn(vals, valid)],
                                [(0, "sum", 0), (0, "count_valid", 0), (0, "min", 0), (0, "max", 0)], sort=False)
        pk = ko[0][0]
        owner = np.mod(pk, world)
        order = np.argsort(owner, kind="stable")
        send_counts = [int(np.count_nonzero(owner == r)) for r in range(world)]
        bufs, flags = [torch.from_numpy(pk[order].copy()), None], [0]
        for data, v, has_mask in ao:
            bufs.append(torch.from_numpy(data[order].copy()))
            bufs.append(torch.from_numpy(v[order].astype(np.uint8)) if has_mask else None)
            flags.append(1 if has_mask else 0)
        comm = D.TorchComm()
        assert comm.world == world and comm.rank == rank
        meta = comm.gather_counts(send_counts + flags, None)
        assert meta.shape == (world, world + len(flags))
        assert [int(x) for x in meta[rank][:world]] == send_counts
        any_valid = meta[:, world:].sum(dim=0) > 0
        recv_counts = [int(meta[s][rank]) for s in range(world)]
        n_local = len(pk)
        for i in range(len(flags)):
            if bufs[2 * i + 1] is None and bool(any_valid[i]):
                bufs[2 * i + 1] = torch.ones(n_local, dtype=torch.uint8)
        got = comm.all_to_all(bufs, send_counts, recv_counts)
        assert all((g is None) == (b is None) for g, b in zip(got, bufs))
        rk = got[0].numpy()
        assert len(rk) == sum(recv_counts)
        assert np.all(np.mod(rk, world) == rank), "received a key this rank does not own"
        # merge (oracle as the stand-in for the CUDA merge): SUM of sums, SUM of counts, MIN of mins, MAX of maxes
        def col(i):
            v = got[2 * i + 1]
            return orc.HostColumn(got[2 * i].numpy(), None if v is None else v.numpy().astype(bool))
        Gm, km, am = orc.groupby([orc.HostColumn(rk)], [col(1), col(2), col(3), col(4)],
                                 [(0, "sum", 0), (1, "sum", 0), (2, "min", 0), (3, "max", 0)])
        np.savez(os.path.join(outdir, f"r{rank}.npz"), keys=km[0][0], s=am[0][0], sv=am[0][1], c=am[1][0], mn=am[2][0], mnv=am[2][1],
                 mx=am[3][0], mxv=am[3][1])
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_exchange_world2_gloo(tmp_path):
    import torch.multiprocessing as mp
    from oracle import oracle as orc
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    parts = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    # expected: the oracle on the concatenated shards
    shards = [_shard(r) for r in range(world)]
    keys = np.concatenate([s[0] for s in shards])
    vals = np.concatenate([s[1] for s in shards])
    valid = np.concatenate([s[2] if s[2] is not None else np.ones(len(s[0]), bool) for s in shards])
    G, ko, ao = orc.groupby([orc.HostColumn(keys)], [orc.HostColumn(vals, valid)],
                            [(0, "sum", 0), (0, "count_valid", 0), (0, "min", 0), (0, "max", 0)])
    got_keys = np.concatenate([p["keys"] for p in parts])
    order = np.argsort(got_keys, kind="stable")
    assert len(got_keys) == G
    np.testing.assert_array_equal(got_keys[order], ko[0][0])
    for name, vname, idx in (("s", "sv", 0), ("mn", "mnv", 2), ("mx", "mxv", 3)):
        data = np.concatenate([p[name] for p in parts])[order]
        v = np.concatenate([p[vname] for p in parts])[order]
        np.testing.assert_array_equal(v, ao[idx][1])
        np.testing.assert_array_equal(data[v], ao[idx][0][ao[idx][1]])
    np.testing.assert_array_equal(np.concatenate([p["c"] for p in parts])[order], ao[1][0].astype(np.int64))
    # every key is owned by exactly one rank
    assert len(np.intersect1d(parts[0]["keys"], parts[1]["keys"])) == 0
