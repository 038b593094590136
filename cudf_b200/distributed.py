"""Multi-GPU hash groupby: one process per GPU, local pre-aggregation -> owner-rank partition of the PARTIALS ->
all-to-all over NCCL (NVLink 5 / NVSwitch) -> merge on the owner.

Rows never move: every rank aggregates the rows it holds with the single-GPU path (GroupBy.aggregate), so the exchange
carries one fixed-width record per (rank, group) instead of one per row. Rank r ends up owning the groups whose key has
hash2(key) % world == r; concatenating the ranks' results gives the groupby of the concatenated input.

Reference precedents: cudf_polars' piecewise-aggregate -> shuffle -> reduce decomposition
(/root/reference/python/cudf_polars/cudf_polars/streaming/groupby.py:322-435: sum -> sum/sum, count -> count/sum,
mean -> (sum, count)/(sum, sum)/div), the shuffle unit cudf::hash_partition
(/root/reference/cpp/libcudf_streaming/src/partition_utils.cpp:56-67,95-116) and streaming_groupby::merge
(/root/reference/cpp/src/groupby/streaming_groupby/merge.cu:30-144).

The collective is torch.distributed (NCCL on GPUs; gloo for the CPU tests of this host logic). Every compute step --
local aggregation, partition, validity repacking, merge, MEAN/VAR/STD finalisation -- is a kernel of libbgb.so reached
through the C ABI (include/bgb.h); nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import torch

from . import _lib
from . import aggregation as agg
from .aggregation import Kind, NullPolicy
from .column import Column, Table, torch_dtype

# ---------------------------------------------------------------------------------------------------------------------
# 1. decomposition of the requested aggregations into partial -> merge -> finalise


_MERGE_KIND = {Kind.SUM: Kind.SUM, Kind.PRODUCT: Kind.PRODUCT, Kind.MIN: Kind.MIN, Kind.MAX: Kind.MAX,
               Kind.SUM_OF_SQUARES: Kind.SUM, Kind.COUNT_VALID: Kind.MERGE_COUNT, Kind.COUNT_ALL: Kind.MERGE_COUNT,
               Kind.MERGE_COUNT: Kind.MERGE_COUNT}  # (merging merged partials: the sum of counts again)


@dataclass
class RequestPlan:
    partial_kinds: list = field(default_factory=list)   # Kind of each partial column computed locally (deduplicated)
    recipes: list = field(default_factory=list)         # one per requested aggregation, see decompose()

    def partial(self, kind: int) -> int:
        if kind not in self.partial_kinds:
            self.partial_kinds.append(kind)
        return self.partial_kinds.index(kind)


def decompose(aggregations) -> RequestPlan:
    """[Aggregation] of one request -> RequestPlan.

    recipes[i] is ("copy", p) when the answer is merged partial column p, ("mean", p_sum, p_count), or
    ("moment", kind, ddof, p_sum, p_sumsq, p_count) for M2 / VARIANCE / STD (sum-of-squares formulation, the one the
    single-GPU finaliser uses, so integer inputs give bit-identical results on any number of ranks)."""
    rp = RequestPlan()
    for a in aggregations:
        k = a.kind
        if k in (Kind.SUM, Kind.PRODUCT, Kind.MIN, Kind.MAX, Kind.SUM_OF_SQUARES, Kind.COUNT_VALID, Kind.COUNT_ALL, Kind.MERGE_COUNT):
            rp.recipes.append(("copy", rp.partial(k)))
        elif k == Kind.MEAN:
            rp.recipes.append(("mean", rp.partial(Kind.SUM), rp.partial(Kind.COUNT_VALID)))
        elif k in (Kind.M2, Kind.VARIANCE, Kind.STD):
            rp.recipes.append(("moment", k, a.ddof, rp.partial(Kind.SUM), rp.partial(Kind.SUM_OF_SQUARES),
                               rp.partial(Kind.COUNT_VALID)))
        else:
            raise _lib.UnsupportedError(2, f"aggregation kind {k} has no distributed decomposition (ARGMIN/ARGMAX refer to "
                                           "local row indices; quantile-type aggregations are not hash aggregations)")
    return rp


def merge_kind(partial_kind: int) -> int:
    return _MERGE_KIND[partial_kind]


# ---------------------------------------------------------------------------------------------------------------------
# 2. the exchange (device-agnostic: NCCL tensors on GPUs, gloo tensors in the CPU tests)


_WIRE_DTYPE = {1: torch.uint8, 2: torch.int16, 4: torch.int32, 8: torch.int64}


def _wire(t: torch.Tensor) -> torch.Tensor:
    """Reinterpret as a dtype every backend can move (uint16/32/64 and bool are not NCCL types)."""
    return t.view(_WIRE_DTYPE[t.element_size()])


def exchange_counts(send_counts, group=None, device=None) -> torch.Tensor:
    """All ranks learn the whole world x len matrix M[src][j] (for j < world: records src sends to rank j) with one
    all-gather; returned on the host."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    mine = torch.tensor(list(send_counts), dtype=torch.int64, device=device)
    allm = torch.empty(world * mine.numel(), dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(allm, mine, group=group)
    return allm.view(world, mine.numel()).cpu()


def exchange_columns(buffers, send_counts, recv_counts, group=None):
    """buffers: [tensor | None] each laid out as `world` contiguous runs of send_counts[r] elements (already permuted by
    owner rank). Returns the received tensors, run s holding what rank s sent here. One all-to-all-v per buffer."""
    import torch.distributed as dist
    total = int(sum(recv_counts))
    out = []
    send_counts = [int(x) for x in send_counts]
    recv_counts = [int(x) for x in recv_counts]
    for b in buffers:
        if b is None:
            out.append(None)
            continue
        src = _wire(b)
        dst = torch.empty(total, dtype=src.dtype, device=src.device)
        dist.all_to_all_single(dst, src, output_split_sizes=recv_counts, input_split_sizes=send_counts, group=group)
        out.append(dst.view(b.dtype))
    return out


class TorchComm:
    """The communicator DistributedGroupBy talks to: torch.distributed over NCCL (GPUs) or gloo (CPU tests)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)

    def gather_counts(self, vec, device):
        return exchange_counts(vec, self.group, device)

    def all_to_all(self, buffers, send_counts, recv_counts):
        return exchange_columns(buffers, send_counts, recv_counts, self.group)

    def peer_connect(self, region_bytes, device):
        import torch.distributed as dist
        if device.type != "cuda" or dist.get_backend(self.group) != "nccl":
            return None
        return _connect_over_torch(self.group, region_bytes, device)


class SingleComm:
    """world == 1: nothing to exchange (the partition / merge steps still run, which is what the 1-GPU tests exercise)."""
    world, rank = 1, 0

    def gather_counts(self, vec, device):
        return torch.tensor([list(vec)], dtype=torch.int64)

    def all_to_all(self, buffers, send_counts, recv_counts):
        return list(buffers)

    def peer_connect(self, region_bytes, device):
        return PeerContext(0, 1, region_bytes, device)


# ---------------------------------------------------------------------------------------------------------------------
# 2b. peer-memory exchange context (bgb_peer_*): every rank's receive buffer mapped into every other rank


class PeerContext:
    """One rank's exchange buffer plus the mappings of its peers' buffers (include/bgb.h, bgb_peer_create ...)."""

    def __init__(self, rank: int, world: int, region_bytes: int, device):
        self.rank, self.world, self.region_bytes, self.device = rank, world, int(region_bytes), device
        self.handle = C.c_void_p()
        with torch.cuda.device(device):
            _lib.check(_lib.lib().bgb_peer_create(rank, world, self.region_bytes, C.byref(self.handle)))

    def export(self) -> bytes:
        buf = C.create_string_buffer(64)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bgb_peer_export(self.handle, buf))
        return buf.raw

    def attach(self, peer_rank: int, handle: bytes):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bgb_peer_attach(self.handle, peer_rank, C.c_char_p(handle)))

    def attach_local(self, peer_rank: int, other: "PeerContext"):
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().bgb_peer_attach_local(self.handle, peer_rank, other.handle))

    def close(self):
        if self.handle:
            with torch.cuda.device(self.device):
                torch.cuda.synchronize()
                _lib.lib().bgb_peer_destroy(self.handle)
            self.handle = C.c_void_p()


def _connect_over_torch(group, region_bytes: int, device):
    """Collective: create this rank's context, swap the IPC handles, map every peer. None when any rank cannot."""
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    ctx, handle = None, None
    try:
        ctx = PeerContext(rank, world, region_bytes, device)
        handle = ctx.export()
    except _lib.BgbError:
        handle = None
    handles = [None] * world
    dist.all_gather_object(handles, handle, group=group)
    ok = all(h is not None for h in handles)
    if ok:
        try:
            for r, h in enumerate(handles):
                if r != rank:
                    ctx.attach(r, h)
        except _lib.BgbError:
            ok = False
    oks = [None] * world
    dist.all_gather_object(oks, ok, group=group)  # also the barrier that makes every buffer ready before its first use
    if all(oks):
        return ctx
    if ctx is not None and ctx.handle:
        ctx.close()
    return None


# ---------------------------------------------------------------------------------------------------------------------
# 3. device steps through the C ABI


def _stream_ptr(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def partition_by_owner(keys: Table, columns, world: int):
    """Permute `columns` (same row count as `keys`) into `world` runs by owner rank = hash2(key) % world
    (bgb_partition_by_key_hash). Returns ([(data, valid_bytes | None)], offsets[world + 1])."""
    L = _lib.lib()
    kcols = keys.columns()
    n = kcols[0].size
    device = kcols[0].data.device
    ck = (_lib.ColumnView * len(kcols))()
    for i, c in enumerate(kcols):
        ck[i] = c.view()
    out = []
    offsets = (C.c_int32 * (world + 1))()
    with torch.cuda.device(device):
        for lo in range(0, max(len(columns), 1), 16):  # kMaxPartValues columns per launch
            chunk = columns[lo:lo + 16]
            cc = (_lib.ColumnView * max(len(chunk), 1))()
            oc = (_lib.MutableColumn * max(len(chunk), 1))()
            bufs = []
            for i, c in enumerate(chunk):
                cc[i] = c.view()
                data = torch.empty(n, dtype=c.data.dtype, device=device)
                vb = torch.empty(n, dtype=torch.uint8, device=device) if c.mask is not None else None
                bufs.append((data, vb))
                oc[i] = _lib.MutableColumn(c.type_id, n, data.data_ptr() if n else None, vb.data_ptr() if vb is not None and n else None, 0, 0)
            _lib.check(L.bgb_partition_by_key_hash(ck, len(kcols), cc, len(chunk), world, oc, offsets, _stream_ptr(device)))
            out.extend(bufs)
    return out, [int(offsets[i]) for i in range(world + 1)]


def column_from_wire(data: torch.Tensor, valid_bytes, type_id: int, scale: int = 0) -> Column:
    """Received (data, validity bytes) -> Column with a packed cudf bitmask (bgb_bytes_to_bitmask)."""
    n = data.numel()
    if valid_bytes is None:
        return Column(data, None, 0, type_id, scale=scale)
    L = _lib.lib()
    device = data.device
    mask = torch.zeros((n + 31) // 32, dtype=torch.int32, device=device)
    nulls = C.c_int32(0)
    with torch.cuda.device(device):
        _lib.check(L.bgb_bytes_to_bitmask(valid_bytes.data_ptr() if n else None, n, mask.data_ptr() if n else None, C.byref(nulls),
                                          _stream_ptr(device)))
    return Column(data, mask, nulls.value, type_id, scale=scale)


def finalize_moments(kind: int, ddof: int, sum_col: Column, sumsq_col, count_col: Column, with_mask: bool, out_type: int = 10) -> Column:
    """MEAN / M2 / VARIANCE / STD from merged SUM / SUM_OF_SQUARES / COUNT columns (bgb_finalize_moments)."""
    L = _lib.lib()
    n = sum_col.size
    device = sum_col.data.device
    data = torch.empty(n, dtype=torch_dtype(out_type), device=device)
    mask = torch.zeros((n + 31) // 32, dtype=torch.int32, device=device) if with_mask else None
    out = _lib.MutableColumn(out_type, n, data.data_ptr() if n else None, mask.data_ptr() if mask is not None and n else None, 0, 0)
    sv, cv = sum_col.view(), count_col.view()
    qv = sumsq_col.view() if sumsq_col is not None else None
    with torch.cuda.device(device):
        _lib.check(L.bgb_finalize_moments(kind, ddof, C.byref(sv), C.byref(qv) if qv is not None else None, C.byref(cv), C.byref(out),
                                          _stream_ptr(device)))
    return Column(data, mask, out.null_count, out_type)


# ---------------------------------------------------------------------------------------------------------------------
# 4. the operator


class DistributedGroupBy:
    """GroupBy over rows sharded across the ranks of a torch.distributed process group (one process per GPU).

    Same call shape as GroupBy (pylibcudf groupby.pyx:33-214): DistributedGroupBy(keys, null_handling).aggregate(requests)
    -> (keys Table, [Table per request]) holding the groups THIS rank owns. `keys` / request values are this rank's rows."""

    def __init__(self, keys: Table, null_handling: int = NullPolicy.EXCLUDE, *, options: dict | None = None, group=None, comm=None,
                 exchange: str = "auto", peer_region_bytes: int = 64 << 20):
        """exchange: "peer" = fused peer-memory path (P2P stores over NVLink, bgb_groupby_build_peer), "nccl" = partition +
        all-to-all-v + merge through torch.distributed, "auto" = peer when every rank can map every other rank's buffer."""
        self._keys = keys
        self._null_handling = null_handling
        self._options = dict(options or {})
        if comm is None:
            import torch.distributed as dist
            comm = TorchComm(group) if dist.is_available() and dist.is_initialized() else SingleComm()
        self._comm = comm
        if exchange not in ("auto", "peer", "nccl"):
            raise ValueError("exchange must be 'auto', 'peer' or 'nccl'")
        self._exchange = exchange
        self._peer_region_bytes = int(peer_region_bytes)
        self.last_exchange = None
        self.last_path = None
        self._local_profile = []
        self._phase_events = []

    # -- profile: the local kernels' device times plus one entry per distributed phase
    @property
    def last_profile(self):
        out = list(self._local_profile)
        for name, e0, e1 in self._phase_events:
            e1.synchronize()
            out.append((name, e0.elapsed_time(e1)))
        return out

    def _phase(self, name, device):
        if not self._options.get("profile"):
            return None
        e0 = torch.cuda.Event(enable_timing=True)
        e0.record(torch.cuda.current_stream(device))
        return (name, e0)

    def _phase_end(self, tok, device):
        if tok is None:
            return
        e1 = torch.cuda.Event(enable_timing=True)
        e1.record(torch.cuda.current_stream(device))
        self._phase_events.append((tok[0], tok[1], e1))

    # -- peer contexts are expensive (cudaMalloc + IPC mapping): one per communicator, kept until it has to grow
    def _peer_context(self, device, grow=False):
        comm = self._comm
        ctx = getattr(comm, "_bgb_peer_ctx", None)
        if ctx is not None and not grow and ctx.region_bytes >= self._peer_region_bytes:
            return ctx
        want = self._peer_region_bytes
        if ctx is not None:
            want = max(want, ctx.region_bytes * (4 if grow else 1))
            ctx.close()
            comm._bgb_peer_ctx = None
        if getattr(comm, "_bgb_peer_failed", False):
            return None
        ctx = comm.peer_connect(want, device)
        if ctx is None:
            comm._bgb_peer_failed = True  # every rank got the same answer (collective): do not try again
        comm._bgb_peer_ctx = ctx
        return ctx

    def _aggregate_peer(self, requests, device):
        from .groupby import GroupBy
        for r in requests:
            decompose(r._aggregations)  # raises UnsupportedError for kinds without a cross-rank merge, before any exchange
        grow = False
        for _ in range(6):
            ctx = self._peer_context(device, grow)
            if ctx is None:
                return None
            gb = GroupBy(self._keys, self._null_handling, options=self._options)
            gb._peer = ctx
            try:
                out = gb.aggregate(requests)
            except _lib.PeerOverflow:
                grow = True  # reported by EVERY rank in the same step, so all ranks take this branch together
                continue
            self._local_profile = list(gb.last_profile)
            self.last_path = gb.last_path
            self.last_exchange = "peer"
            return out
        raise _lib.CudaError(3, "peer exchange region could not be grown enough")

    def aggregate(self, requests):
        from .groupby import GroupBy, GroupByRequest
        comm = self._comm
        world = comm.world
        device = self._keys.columns()[0].data.device
        self._phase_events = []
        if self._exchange in ("auto", "peer"):
            out = self._aggregate_peer(requests, device)
            if out is not None:
                return out
            if self._exchange == "peer":
                raise _lib.UnsupportedError(2, "peer-memory exchange is not available between these ranks")
        self.last_exchange = "nccl"

        # (1) local pre-aggregation of this rank's rows: one partial column per (request, partial kind)
        plans = [decompose(r._aggregations) for r in requests]
        local_reqs = [GroupByRequest(r._values, [agg.Aggregation(k) for k in p.partial_kinds]) for r, p in zip(requests, plans)]
        local = GroupBy(self._keys, self._null_handling, options=self._options)
        pkeys, presults = local.aggregate(local_reqs)
        self._local_profile = list(local.last_profile)
        self.last_path = local.last_path
        nk = pkeys.num_columns()
        flat = list(pkeys.columns()) + [c for t in presults for c in t.columns()]

        # (2) owner-rank partition of the partial table (keys travel with the partial columns)
        tok = self._phase("dist_partition", device)
        bufs, offsets = partition_by_owner(pkeys, flat, world)
        self._phase_end(tok, device)
        send_counts = [offsets[r + 1] - offsets[r] for r in range(world)]

        # (3) exchange: the count matrix (+ which columns carry validity anywhere), then one all-to-all-v per buffer
        tok = self._phase("dist_exchange", device)
        flags = [1 if vb is not None else 0 for _, vb in bufs]
        meta = comm.gather_counts(send_counts + flags, device)
        any_valid = meta[:, world:].sum(dim=0) > 0
        recv_counts = [int(meta[s][comm.rank]) for s in range(world)]
        if sum(recv_counts) > 2**31 - 1:
            raise _lib.LogicError(1, "a rank would own more than 2^31-1 partial groups (cudf::size_type)")
        n_local = offsets[world]
        wire = []
        for i, (data, vb) in enumerate(bufs):
            if vb is None and bool(any_valid[i]):
                vb = torch.ones(n_local, dtype=torch.uint8, device=device)  # another rank's partial has nulls here
            wire.extend([data, vb])
        got = comm.all_to_all(wire, send_counts, recv_counts)
        self._phase_end(tok, device)

        # (4) merge on the owner: groupby of the received partials with the merge kinds
        tok = self._phase("dist_merge", device)
        cols = [column_from_wire(got[2 * i], got[2 * i + 1], flat[i].type_id, flat[i].scale) for i in range(len(flat))]
        mkeys = Table(cols[:nk])
        merge_reqs, k = [], nk
        for p in plans:
            for kind in p.partial_kinds:
                merge_reqs.append(GroupByRequest(cols[k], [agg.Aggregation(merge_kind(kind))]))
                k += 1
        merge_opts = {kk: v for kk, v in self._options.items() if kk in ("profile",)}
        merger = GroupBy(mkeys, self._null_handling, options=merge_opts)
        okeys, merged = merger.aggregate(merge_reqs)
        self.last_merge_path = merger.last_path

        # (5) finalise: copy-through or MEAN / M2 / VARIANCE / STD from the merged SUM / SUMSQ / COUNT columns
        results = self._finalize(requests, plans, merged, device)
        self._phase_end(tok, device)
        return okeys, results


    # -- several shards per rank (strong scaling: BASELINE configs[4] holds 8 shards whatever the number of GPUs)
    def aggregate_shards(self, shards):
        """shards: [(keys Table, [GroupByRequest])] -- the pieces of this rank's rows, each at most 2^31-1 rows (cudf::size_type),
        all with the same column types and aggregations. Every shard is pre-aggregated with the single-GPU path into PARTIAL
        columns (decompose()), the shards' partial tables are folded together on this rank with the merge kinds (in batches
        of ~4.5e8 pending rows: options["fold_rows"]), and the folded table goes through the ordinary distributed
        merge (peer-memory or NCCL exchange by owner rank). One shard is exactly aggregate(). The decomposition is the one
        cudf_polars uses across partitions (/root/reference/python/cudf_polars/cudf_polars/streaming/groupby.py:322-435); the
        fold is streaming_groupby::merge (/root/reference/cpp/src/groupby/streaming_groupby/merge.cu:30-144)."""
        from .groupby import GroupBy, GroupByRequest
        it = iter(shards)  # consumed lazily: a caller may stage shard i + 1 (host -> device) while shard i is aggregated
        first = next(it)
        second = next(it, None)
        if second is None:
            self._keys = first[0]
            return self.aggregate(first[1])
        requests0 = first[1]
        plans = [decompose(r._aggregations) for r in requests0]
        device = first[0].columns()[0].data.device
        self._phase_events = []
        self._local_profile = []
        nk = first[0].num_columns()

        def chain():
            yield first
            yield second
            yield from it
        merge_opts = {kk: v for kk, v in self._options.items() if kk in ("profile",)}

        def merge_requests(cols):
            reqs, k = [], nk
            for p in plans:
                for kind in p.partial_kinds:
                    reqs.append(GroupByRequest(cols[k], [agg.Aggregation(merge_kind(kind))]))
                    k += 1
            return reqs

        def concat(tables):
            """partial tables (lists of Columns with offset 0) -> one, ONE torch.cat per column (no pairwise intermediates: every
            intermediate would stay in torch's block cache, out of reach of the library's pool); validity through bytes when any
            table carries a mask"""
            out = []
            for ci in range(len(tables[0])):
                cs = [t[ci] for t in tables]
                data = torch.cat([c.data[:c.size] for c in cs])
                if all(c.mask is None for c in cs):
                    out.append(Column(data, None, 0, cs[0].type_id, scale=cs[0].scale))
                    continue
                vb = torch.cat([_valid_bytes(c) for c in cs])
                out.append(column_from_wire(data, vb, cs[0].type_id, cs[0].scale))
            return out

        def release_caches():
            """memory pressure: torch's and the library's cached blocks go back to the driver (each allocator then sees the other's)"""
            torch.cuda.synchronize(device)
            torch.cuda.empty_cache()
            _lib.lib().bgb_release_cached_memory()

        def fold(tables):
            """several partial tables -> one (concatenate, then ONE merge groupby)"""
            try:
                cols = concat(tables)
            except torch.OutOfMemoryError:
                release_caches()
                cols = concat(tables)
            tables.clear()  # the concatenation replaces them
            reqs_m = merge_requests(cols)
            try:
                fk, fr = GroupBy(Table(cols[:nk]), self._null_handling, options=merge_opts).aggregate(reqs_m)
            except (_lib.CudaError, torch.OutOfMemoryError) as e:
                if isinstance(e, _lib.CudaError) and e.status != 5:  # BGB_OUT_OF_MEMORY: once more with both caches emptied
                    raise
                release_caches()
                fk, fr = GroupBy(Table(cols[:nk]), self._null_handling, options=merge_opts).aggregate(reqs_m)
            return list(fk.columns()) + [t.columns()[0] for t in fr]

        # partial tables wait for each other until ~fold_rows rows are pending, then one fold replaces them (one merge groupby over
        # 4e8 rows costs about as much as one over 2e8, so folding pairwise would pay it once per shard)
        fold_rows = int(self._options.get("fold_rows", 450_000_000))
        pending, pending_rows = [], 0
        for keys, reqs in chain():
            local_reqs = [GroupByRequest(r._values, [agg.Aggregation(k) for k in p.partial_kinds]) for r, p in zip(reqs, plans)]
            tok = self._phase("dist_local", device)
            local = GroupBy(keys, self._null_handling, options=self._options)
            try:
                pkeys, presults = local.aggregate(local_reqs)
            except (_lib.CudaError, torch.OutOfMemoryError) as e:
                if isinstance(e, _lib.CudaError) and e.status != 5:
                    raise
                release_caches()
                pkeys, presults = local.aggregate(local_reqs)
            self._phase_end(tok, device)
            self._local_profile.extend(local.last_profile)
            self.last_path = local.last_path
            flat = list(pkeys.columns()) + [c for t in presults for c in t.columns()]
            pending.append(flat)
            pending_rows += flat[0].size
            del pkeys, presults, flat
            if len(pending) > 1 and pending_rows >= fold_rows:
                tok = self._phase("dist_fold", device)
                pending = [fold(pending)]
                pending_rows = pending[0][0].size
                self._phase_end(tok, device)
        if len(pending) > 1 and self._comm.world == 1:
            tok = self._phase("dist_fold", device)
            pending = [fold(pending)]
            self._phase_end(tok, device)
        if len(pending) > 1:
            # several ranks: the exchange step's local build IS a fold (a groupby of partial rows with the merge kinds), so the
            # pending tables go into it concatenated instead of being folded first and re-aggregated there (one pass less)
            try:
                running = concat(pending)
            except torch.OutOfMemoryError:
                release_caches()
                running = concat(pending)
            pending.clear()
        else:
            running = pending[0]

        if self._comm.world == 1:
            # one rank: the folded table already holds every group once -- nothing to exchange, nothing to merge
            self.last_exchange = "none"
            return Table(running[:nk]), self._finalize(requests0, plans, [Table([c]) for c in running[nk:]], device)
        # the folded partial table through the distributed merge (its requests are merge kinds of single partial columns)
        dg = DistributedGroupBy(Table(running[:nk]), self._null_handling, options=merge_opts, comm=self._comm, exchange=self._exchange,
                                peer_region_bytes=self._peer_region_bytes)
        okeys, merged = dg.aggregate(merge_requests(running))
        self.last_exchange = dg.last_exchange
        self._local_profile.extend(dg._local_profile)
        self._phase_events.extend(dg._phase_events)
        return okeys, self._finalize(requests0, plans, merged, device)

    def _finalize(self, requests, plans, merged, device):
        """merged: one single-column Table per partial column, request-major -> one result Table per request"""
        results, k = [], 0
        for r, p in zip(requests, plans):
            part = [merged[k + j].columns()[0] for j in range(len(p.partial_kinds))]
            k += len(p.partial_kinds)
            outs = []
            for rec in p.recipes:
                if rec[0] == "copy":
                    outs.append(part[rec[1]])
                elif rec[0] == "mean":
                    s_, c_ = part[rec[1]], part[rec[2]]
                    out_type = _lib.lib().bgb_target_type(r._values.type_id, Kind.MEAN)
                    outs.append(finalize_moments(Kind.MEAN, 0, s_, None, c_, s_.mask is not None, out_type))
                else:
                    _, kind, ddof, si, qi, ci = rec
                    outs.append(finalize_moments(kind, ddof, part[si], part[qi], part[ci], kind != Kind.M2))
            results.append(Table(outs))
        return results


def _valid_bytes(col: Column) -> torch.Tensor:
    """validity of a Column (offset 0) as one byte per row (1 = valid)"""
    n = col.size
    if col.mask is None:
        return torch.ones(n, dtype=torch.uint8, device=col.data.device)
    words = col.mask.view(torch.int32)
    idx = torch.arange(n, device=col.data.device, dtype=torch.int64)
    return ((words[idx >> 5] >> (idx & 31).to(torch.int32)) & 1).to(torch.uint8)
