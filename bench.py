#!/usr/bin/env python
"""bench.py -- hash groupby-aggregate throughput on B200 (rows/s + fraction of the HBM roofline).

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.
A "step" is one full groupby call (build + emit: hash, aggregate, compact, gather/finalise) over one batch of synthetic
columns that are already resident in HBM; `e2e` is the same call through the C ABI's host-buffer entry point
(bgb_groupby_host) with pinned HOST columns, H2D/D2H inside the timed region.

Workloads (BASELINE.json configs / SURVEY.md 8d):
  A9g    int64 key x%1e8, int64 value, SUM+COUNT, N=1e9     <- HEADLINE at --gpus 1: the largest single-GPU configuration, the
                                                               shape north_star's target is quoted on (1e9 rows, 1e8 groups)
  A9     as A9g with 1e4 groups                              (north_star's other target shape)
  A      as A9 with N=1e8                                    (configs[1])
  AS     as A with SPARSE key values (splitmix64 of x%1e4)   (the shared-memory table cannot direct-address these)
  B      int32 key x%1e8, float64 value 10% null, SUM+MIN+MAX, N=1e9   (configs[2])
  Q1     2 int8 keys (4 groups), 2 float64 values, MEAN+VAR+COUNT each, N=6e8   (configs[3])
  C1     2 int64 keys 32x32, 2 float64 MEANs, N=1e7          (configs[0]'s shape on the GPU)
  MG     configs[4]: 8 shards x 1e9 rows (shard s = seed 42+s), int64 key x%1e8, SUM+COUNT -- STRONG scaling: the 8 shards
         are spread over the N ranks (8/N each; at N=1 every shard is fed as two 5e8-row halves so that inputs +
         scratch fit the GPU), per-shard partials are folded on the rank, the folded table is exchanged by owner rank (peer-memory
         stores over NVLink, or --exchange nccl for the all-to-all) and merged. This is what --gpus N>1 runs.
With no --workload, --gpus 1 times A9g as the headline and then A9, B, Q1, C1, A, AS as a "sweep" array in the same JSON line
(the reference's own precedent for one run across cardinalities: cpp/benchmarks/groupby/group_max.cpp:79-165).
Every workload's result is VERIFIED outside the timed region: per-group against the CPU oracle on a prefix of the rows
(>= 5e7 where the workload has them) and through sum / count / unique-key invariants on the full result.

`--impl reference` times the CPU restatement of the reference algorithm (oracle/, all host threads) on a bounded sample
of the same workload: libcudf itself cannot be built offline, so the oracle port is the reference arm.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "groupby_agg_rows_per_sec"
UNIT = "rows/s"

WORKLOADS = {
    # name: (rows, groups, dtype, description)
    "A9g": (1_000_000_000, 100_000_000, "int64", "single int64 key SUM/COUNT, 1e9 rows, 1e8 groups (north_star target shape; largest single-GPU configuration)"),
    "A9": (1_000_000_000, 10_000, "int64", "single int64 key SUM/COUNT, 1e9 rows, 1e4 groups (north_star target shape)"),
    "A": (100_000_000, 10_000, "int64", "single int64 key SUM/COUNT, 100M rows, 1e4 groups (BASELINE.json configs[1])"),
    "AS": (100_000_000, 10_000, "int64", "as A with sparse key values: key = splitmix64(x % 1e4) (hashed shared-memory regime)"),
    "B": (1_000_000_000, 100_000_000, "f64", "int32 key SUM/MIN/MAX float64 with 10% nulls, 1e9 rows, 1e8 groups (configs[2])"),
    "Q1": (600_000_000, 4, "f64", "TPC-H Q1-style two int8 keys MEAN/VAR/COUNT over 6e8 rows (configs[3])"),
    "C1": (10_000_000, 1024, "f64", "df.groupby(['A','B']).mean(): two int64 keys in [0,32), two float64 values, 10M rows (configs[0] shape on the GPU)"),
    "MG": (8_000_000_000, 100_000_000, "int64", "int64 key SUM/COUNT, 8e9 rows as 8 shards x 1e9, 1e8 groups, partials exchanged by owner rank (configs[4])"),
}
SWEEP = ["A9", "B", "Q1", "C1", "A", "AS"]
SPARSE_SEED = 7_777


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------------ synthetic inputs
def algorithmic_bytes(workload: str, n: int, g: int) -> float:
    """minimum reads + minimum writes of one call."""
    if workload in ("A", "A9", "A9g", "AS", "MG"):
        return 16.0 * n + 20.0 * g
    if workload == "B":
        return 12.125 * n + (4 + 24 + 3 / 8) * g
    if workload == "Q1":
        return 18.0 * n
    if workload == "C1":
        return 32.0 * n + 32.0 * g
    raise ValueError(workload)


def _torch_mix64(z):
    """SplitMix64 finaliser on int64 tensors (wrapping arithmetic; logical shifts through masks)."""
    import torch
    def lsr(x, k):
        return (x >> k) & ((1 << (64 - k)) - 1)
    c1 = torch.tensor(-4658895280553007687, dtype=torch.int64, device=z.device)   # 0xBF58476D1CE4E5B9
    c2 = torch.tensor(-7723592293110705685, dtype=torch.int64, device=z.device)   # 0x94D049BB133111EB
    z = (z ^ lsr(z, 30)) * c1
    z = (z ^ lsr(z, 27)) * c2
    return z ^ lsr(z, 31)


def make_device_inputs(workload: str, n: int, g: int, seed: int, device, first_row: int = 0):
    """Columns generated ON the device by the library's counter-based SplitMix64 generators (bit-identical to the oracle's)."""
    import ctypes as C

    import torch

    import cudf_b200 as cb
    from cudf_b200 import _lib, aggregation as A
    L = _lib.lib()
    s = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)

    def gen_int(dtype, tid, sd, mod, add=0):
        t = torch.empty(n, dtype=dtype, device=device)
        _lib.check(L.bgb_gen_uniform_int(C.c_void_p(t.data_ptr()), tid, n, sd, first_row, mod, add, s))
        return t

    def gen_real(sd, mod, add=0.0):
        t = torch.empty(n, dtype=torch.float64, device=device)
        _lib.check(L.bgb_gen_uniform_real(C.c_void_p(t.data_ptr()), cb.TypeId.FLOAT64, n, sd, first_row, mod, add, s))
        return t

    if workload in ("A", "A9", "A9g", "AS", "MG"):
        kt = gen_int(torch.int64, cb.TypeId.INT64, seed, g)
        if workload == "AS":
            step = 1 << 26
            for lo in range(0, n, step):  # in place, chunked: the mix needs a few temporaries
                kt[lo:lo + step] = _torch_mix64(kt[lo:lo + step] + (SPARSE_SEED * 0x1000003 + 0x9E3779B97F4A7C15 - (1 << 64)))
        k = cb.Column(kt)
        v = cb.Column(gen_int(torch.int64, cb.TypeId.INT64, seed + 1, 2001, -1000))
        return cb.Table([k]), [cb.GroupByRequest(v, [A.sum(), A.count()])]
    if workload == "B":
        k = cb.Column(gen_int(torch.int32, cb.TypeId.INT32, seed, g))
        vd = gen_real(seed + 1, 1_000_000)
        mask = torch.empty((n + 31) // 32, dtype=torch.int32, device=device)
        nc = C.c_int32(0)
        _lib.check(L.bgb_gen_null_mask(C.c_void_p(mask.data_ptr()), n, seed + 2, first_row, 10, C.byref(nc), s))
        v = cb.Column(vd, mask, nc.value)
        return cb.Table([k]), [cb.GroupByRequest(v, [A.sum(), A.min(), A.max()])]
    if workload == "C1":
        ka = cb.Column(gen_int(torch.int64, cb.TypeId.INT64, seed, 32))
        kb = cb.Column(gen_int(torch.int64, cb.TypeId.INT64, seed + 1, 32))
        c = cb.Column(gen_real(seed + 2, 1_000_000))
        d = cb.Column(gen_real(seed + 3, 1_000_000))
        return cb.Table([ka, kb]), [cb.GroupByRequest(c, [A.mean()]), cb.GroupByRequest(d, [A.mean()])]
    if workload == "Q1":
        # (returnflag, linestatus) joint weights (A,F) 24.7%, (N,F) 0.6%, (N,O) 50%, (R,F) 24.7%: one categorical draw
        # picks the pair, so both key columns come from the same x(seed,row)
        cum = (C.c_int32 * 4)(int(0.247 * 65536), int(0.253 * 65536), int(0.753 * 65536), 65536)
        rf = torch.empty(n, dtype=torch.int8, device=device)
        ls = torch.empty(n, dtype=torch.int8, device=device)
        _lib.check(L.bgb_gen_categorical_i8(C.c_void_p(rf.data_ptr()), n, seed, first_row, cum, bytes([0, 1, 1, 2]), 4, s))
        _lib.check(L.bgb_gen_categorical_i8(C.c_void_p(ls.data_ptr()), n, seed, first_row, cum, bytes([0, 0, 1, 0]), 4, s))
        q = cb.Column(gen_real(seed + 1, 50, 1.0))
        p = cb.Column(gen_real(seed + 2, 104_100, 900.0))
        aggs = lambda: [A.mean(), A.variance(1), A.count()]  # noqa: E731
        return cb.Table([cb.Column(rf), cb.Column(ls)]), [cb.GroupByRequest(q, aggs()), cb.GroupByRequest(p, aggs())]
    raise ValueError(workload)


def make_host_inputs_numpy(workload: str, n: int, g: int, seed: int):
    """The same columns on the host (numpy), for the CPU legs and the checker. Identical values to make_device_inputs."""
    from oracle import oracle as orc
    rows = np.arange(n, dtype=np.uint64)
    H = orc.HostColumn
    if workload in ("A", "A9", "A9g", "AS", "MG"):
        k = orc.splitmix64(seed, rows) % np.uint64(g)
        if workload == "AS":
            k = orc.splitmix64(SPARSE_SEED, k)
        v = (orc.splitmix64(seed + 1, rows) % np.uint64(2001)).astype(np.int64) - 1000
        return [H(k.astype(np.int64))], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)]
    if workload == "B":
        k = (orc.splitmix64(seed, rows) % np.uint64(g)).astype(np.int32)
        v = (orc.splitmix64(seed + 1, rows) % np.uint64(1_000_000)).astype(np.float64)
        valid = (orc.splitmix64(seed + 2, rows) % np.uint64(10)) != 0
        return [H(k)], [H(v, valid)], [(0, "sum", 1), (0, "min", 1), (0, "max", 1)]
    if workload == "C1":
        ka = (orc.splitmix64(seed, rows) % np.uint64(32)).astype(np.int64)
        kb = (orc.splitmix64(seed + 1, rows) % np.uint64(32)).astype(np.int64)
        c = (orc.splitmix64(seed + 2, rows) % np.uint64(1_000_000)).astype(np.float64)
        d = (orc.splitmix64(seed + 3, rows) % np.uint64(1_000_000)).astype(np.float64)
        return [H(ka), H(kb)], [H(c), H(d)], [(0, "mean", 1), (1, "mean", 1)]
    if workload == "Q1":
        x = (orc.splitmix64(seed, rows) & np.uint64(0xFFFF)).astype(np.int64)
        cum = [int(0.247 * 65536), int(0.253 * 65536), int(0.753 * 65536), 65536]
        cat = np.searchsorted(np.array(cum), x, side="right").clip(0, 3)
        rf = np.array([0, 1, 1, 2], dtype=np.int8)[cat]
        ls = np.array([0, 0, 1, 0], dtype=np.int8)[cat]
        q = (orc.splitmix64(seed + 1, rows) % np.uint64(50)).astype(np.float64) + 1.0
        p = (orc.splitmix64(seed + 2, rows) % np.uint64(104_100)).astype(np.float64) + 900.0
        aggs = [(0, "mean", 1), (0, "variance", 1), (0, "count_valid", 1), (1, "mean", 1), (1, "variance", 1), (1, "count_valid", 1)]
        return [H(rf), H(ls)], [H(q), H(p)], aggs
    raise ValueError(workload)


# ------------------------------------------------------------------------------------------------ CPU legs
def cpu_leg(workload: str, n_full: int, g: int, seed: int, sample_rows: int, repeats: int = 1):
    """Times the oracle (CPU restatement of the reference algorithm, OpenMP over all host cores) on the first
    `sample_rows` rows of the workload. Returns (rows/s, cores, sample description, seconds per pass)."""
    from oracle import oracle as orc
    n = min(n_full, sample_rows)
    keys, vals, aggs = make_host_inputs_numpy(workload, n, g, seed)
    cores = os.cpu_count() or 1
    best = None
    for _ in range(max(1, repeats)):
        t0 = time.perf_counter()
        orc.groupby(keys, vals, aggs, threads=cores, sort=False)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return n / best, cores, f"first {n} rows of the {n_full}-row workload, {g} key values, oracle/groupby_oracle.c -O3 -fopenmp", best


def pandas_leg(workload: str, n_full: int, g: int, seed: int, sample_rows: int):
    """pandas groupby (the cudf.pandas CPU fallback path), single-threaded, same sample."""
    try:
        import pandas as pd
    except Exception:
        return None
    n = min(n_full, sample_rows)
    keys, vals, aggs = make_host_inputs_numpy(workload, n, g, seed)
    cols = {f"k{i}": c.data for i, c in enumerate(keys)}
    for i, c in enumerate(vals):
        d = c.data.astype(np.float64) if c.valid is not None else c.data
        if c.valid is not None:
            d = d.copy()
            d[~c.valid] = np.nan
        cols[f"v{i}"] = d
    df = pd.DataFrame(cols)
    spec = {}
    for vi, kind, _ in aggs:
        spec.setdefault(f"v{vi}", []).append({"count_valid": "count", "variance": "var"}.get(kind, kind))
    t0 = time.perf_counter()
    df.groupby([f"k{i}" for i in range(len(keys))], sort=False).agg(spec)
    dt = time.perf_counter() - t0
    return {"value": n / dt, "unit": UNIT, "cores": 1, "rows": n, "pandas": pd.__version__}


def default_workload(args):
    if args.workload:
        return args.workload
    return "A9g" if args.gpus <= 1 and int(os.environ.get("WORLD_SIZE", "1")) <= 1 else "MG"


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = default_workload(args)
    n, g, dtype, desc = WORKLOADS[wl]
    n = args.rows or n
    g = args.groups or g
    # each step = a bounded sample of the workload (the whole --steps/--warmup run has to end within a few minutes on 16
    # host cores): the first `sample` rows with the workload's key domain, i.e. the same cardinality regime
    sample = min(n, args.cpu_sample_rows)
    gen_wl = "A9g" if wl == "MG" else wl
    keys, vals, aggs = make_host_inputs_numpy(gen_wl, sample, g, 42)
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    for _ in range(min(args.warmup, 1)):
        orc.groupby(keys, vals, aggs, threads=cores, sort=False)
    steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(steps):
        orc.groupby(keys, vals, aggs, threads=cores, sort=False)
    dt = (time.perf_counter() - t0) / steps
    value = sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 1), "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong" if wl == "MG" else "weak", "vs_baseline": None,
        "dtype": dtype, "data": "synthetic",
        "config": {"workload": f"{wl}: {desc}", "rows": n, "groups": g, "sample_rows_per_step": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"each step = first {sample} rows of the workload (same {g}-value key domain); libcudf cannot be built "
                                   "offline, so the reference arm is the CPU restatement in oracle/ (OpenMP, all host cores)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(line)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock + throttle reasons DURING the timed region, through NVML (pynvml): a 20 ms background thread (its queries
    run with the GIL released; tools/nvml_cost.py), plus one sample right after the last step is enqueued when the region
    was shorter than the period; the max-clock query costs milliseconds and is made once before the region. (An `nvidia-smi -lms` child polling every
    50 ms stretched 3 ms steps to 14 ms on this box -- the driver serialises its queries with kernel launches -- so the
    sampler is kept this light; the samples are still taken while the steps run.)"""

    def __init__(self, index: int):
        self.index = index
        self.samples = []
        self._stop = threading.Event()
        self._t = None
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)  # slow query (ms): once, up front
        except Exception:
            self.nv = None
            self.h = None

    def sample(self):
        if self.h is None:
            return
        try:
            nv = self.nv
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)  # microseconds: cached by the driver
            mx = self.max_mhz
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            self.samples.append((sm, mx, int(r)))
        except Exception:
            pass

    def _run(self):
        while not self._stop.wait(0.02):
            self.sample()

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self):
        self._stop.set()
        if self._t:
            self._t.join(timeout=2)
        names = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        reasons = set()
        for _, _, r in self.samples:
            for bit, nm in names.items():
                if r & bit:
                    reasons.add(nm)
        sm = [x[0] for x in self.samples]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max((x[1] for x in self.samples), default=None),
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self.h is not None else "unavailable"}


# ------------------------------------------------------------------------------------------------ output plumbing
_REAL_STDOUT = None


def _quiet_stdout():
    """Libraries (NCCL's version banner, torchrun) may write to fd 1; the contract is ONE JSON line on stdout. Everything
    else goes to stderr: fd 1 is pointed at fd 2 for the run and the JSON line is written to the saved descriptor."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.dup(1)
        os.dup2(2, 1)


def emit_line(line):
    data = (json.dumps(line) + "\n").encode()
    sys.stdout.flush()
    if _REAL_STDOUT is None:
        os.write(1, data)
    else:
        os.write(_REAL_STDOUT, data)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


# ------------------------------------------------------------------------------------------------ verification (checker)
def _combined_key(cols):
    """key column tensors -> ONE int64 tensor that orders and identifies the groups (bench keys are one int64 / int32 column or
    two small-integer columns)"""
    import torch
    if len(cols) == 1:
        return cols[0].to(torch.int64)
    out = cols[0].to(torch.int64)
    for c in cols[1:]:
        out = out * (1 << 20) + c.to(torch.int64)
    return out


def verify_workload(wl, keys, reqs, n, g, seed, prefix_rows, force_path=0):
    """(a) per-group comparison with the CPU oracle on the first `prefix_rows` rows (the checker, never the thing measured):
    both results are ordered by key ON THE DEVICE (torch.sort) and compared there -- keys, counts, integer sums, MIN / MAX
    bit-exact, float results within 1e-9 relative (north_star's tolerance);
    (b) invariants on the FULL result: sum of counts == valid rows, sum of sums == column total, every key once.
    Returns (ok, details)."""
    import torch

    import cudf_b200 as cb
    from oracle import oracle as orc
    details = {}
    t0 = time.perf_counter()
    dev = keys.columns()[0].data.device
    # ---- (a) prefix against the oracle
    p = min(n, prefix_rows)
    pkeys = cb.Table([c.sliced(0, p) for c in keys.columns()])
    preqs = [cb.GroupByRequest(r._values.sliced(0, p), r._aggregations) for r in reqs]
    kt, rts = cb.GroupBy(pkeys, options={"force_path": force_path}).aggregate(preqs)
    gkey = _combined_key([c.data for c in kt.columns()])
    gorder = torch.argsort(gkey)
    hk, hv, aggs = make_host_inputs_numpy(wl, p, g, seed)
    t1 = time.perf_counter()
    Go, ko, ao = orc.groupby(hk, hv, aggs, threads=os.cpu_count() or 1, sort=False)
    t2 = time.perf_counter()
    ok = kt.num_rows() == Go
    if ok:
        okey = _combined_key([torch.from_numpy(np.ascontiguousarray(d)).to(dev) for d, *_ in ko])
        oorder = torch.argsort(okey)
        ok = bool(torch.equal(gkey[gorder], okey[oorder]))
        flat = [c for t in rts for c in t.columns()]
        for i, c in enumerate(flat):
            od = torch.from_numpy(np.ascontiguousarray(ao[i][0])).to(dev)[oorder]
            ov = torch.from_numpy(np.ascontiguousarray(ao[i][1])).to(dev)[oorder]
            gd = c.data[gorder]
            if c.mask is not None:
                words = c.mask
                idx = gorder.to(torch.int64)
                gv = ((words[idx >> 5] >> (idx & 31).to(torch.int32)) & 1).to(torch.bool)
            else:
                gv = torch.ones(gd.numel(), dtype=torch.bool, device=dev)
            ok = ok and bool(torch.equal(gv, ov))
            if not ok:
                break
            a, b2 = gd[gv], od[ov]
            if a.dtype.is_floating_point:
                ok = ok and bool(torch.allclose(a, b2.to(a.dtype), rtol=1e-9, atol=0.0))
            else:
                ok = ok and a.dtype == b2.dtype and bool(torch.equal(a, b2))
        del flat
    details["oracle_seconds"] = round(t2 - t1, 2)
    details["oracle_prefix_rows"] = p
    details["oracle_prefix_groups"] = int(Go)
    details["oracle_prefix_ok"] = bool(ok)
    del kt, rts, hk, hv, ko, ao
    # ---- (b) full-size invariants
    kt, rts = cb.GroupBy(keys, options={"force_path": force_path}).aggregate(reqs)
    G = kt.num_rows()
    inv = True
    k0 = kt.columns()[0].data
    if kt.num_columns() == 1:
        inv = inv and int(torch.unique(k0).numel()) == G
    else:
        combo = kt.columns()[0].data.to(torch.int64) * 65536 + kt.columns()[1].data.to(torch.int64)
        inv = inv and int(torch.unique(combo).numel()) == G
    for r, t in zip(reqs, rts):
        vcol = r._values
        valid_rows = n - (vcol.null_count if vcol.mask is not None else 0)
        for a, c in zip(r._aggregations, t.columns()):
            if a.kind == cb.aggregation.Kind.COUNT_VALID:
                inv = inv and int(c.data.sum(dtype=torch.int64).item()) == valid_rows
            elif a.kind == cb.aggregation.Kind.SUM:
                if vcol.mask is None:
                    tot = vcol.data.sum(dtype=vcol.data.dtype if vcol.data.dtype.is_floating_point else torch.int64)
                    got = c.data.sum(dtype=c.data.dtype)
                    inv = inv and (tot.item() == got.item())  # int64: exact (wrapping); f64: whole-number inputs, exact below 2^53
    for r, t in zip(reqs, rts):
        kinds = [a.kind for a in r._aggregations]
        if cb.aggregation.Kind.MIN in kinds and cb.aggregation.Kind.MAX in kinds:
            cmin, cmax = t.columns()[kinds.index(cb.aggregation.Kind.MIN)], t.columns()[kinds.index(cb.aggregation.Kind.MAX)]
            both = torch.ones(G, dtype=torch.bool, device=cmin.data.device) if cmin.mask is None else None
            if both is not None:
                inv = inv and bool((cmin.data <= cmax.data).all().item())
            else:
                inv = inv and cmin.null_count == cmax.null_count
    details["full_groups"] = int(G)
    details["full_invariants_ok"] = bool(inv)
    details["seconds"] = round(time.perf_counter() - t0, 2)
    return bool(ok and inv), details


# ------------------------------------------------------------------------------------------------ single-GPU measurement
def measure_single(wl, n, g, seed, device, steps, warmup, force_path, L):
    """Timed region: `steps` calls with kernel profiling OFF (CUDA events around the region on the call's stream); kernel
    times come from a separate profiled pass."""
    import torch

    import cudf_b200 as cb
    keys, reqs = make_device_inputs(wl, n, g, seed, device)
    torch.cuda.synchronize()
    gb = cb.GroupBy(keys, options={"force_path": force_path})
    for _ in range(max(warmup, 3)):
        gb.aggregate(reqs)
    torch.cuda.synchronize()
    launches0 = L.bgb_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    host_ms = []
    e0.record()
    for _ in range(steps):
        th = time.perf_counter()
        gb.aggregate(reqs)
        host_ms.append((time.perf_counter() - th) * 1e3)
    e1.record()
    torch.cuda.synchronize()
    total_ms = e0.elapsed_time(e1)
    launches = L.bgb_kernel_launches() - launches0
    # separate profiled pass: per-kernel device times (events on the launch stream around every kernel)
    gp = cb.GroupBy(keys, options={"force_path": force_path, "profile": 1})
    acc = {}
    for _ in range(3):
        gp.aggregate(reqs)
        for name, ms in gp.last_profile:
            acc.setdefault(name, []).append(ms)
    kern_ms = {k: sum(v) / len(v) for k, v in acc.items()}
    return {"keys": keys, "reqs": reqs, "ms_per_step": total_ms / steps, "path": gb.last_path, "kernels_ms": kern_ms,
            "launches": int(launches), "host_ms": host_ms}


def roofline_of(wl, n, g, m, peak, peak_src):
    kern_ms = m["kernels_ms"]
    dom = max(kern_ms, key=kern_ms.get) if kern_ms else None
    alg = algorithmic_bytes(wl, n, g)
    if not dom:
        return None
    ach = alg / (kern_ms[dom] * 1e-3) / 1e9
    roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
            "peak_source": peak_src, "kernel_ms": kern_ms[dom], "algorithmic_bytes": alg, "traffic": None,
            "frac_of_nominal_8000": ach / 8000.0,
            "whole_call": {"achieved": alg / (m["ms_per_step"] * 1e-3) / 1e9, "frac": alg / (m["ms_per_step"] * 1e-3) / 1e9 / peak},
            "kernels_ms": {k: round(v, 4) for k, v in kern_ms.items()},
            "kernels_note": "per-kernel times from a separate profiled pass (profiling is off in the timed steps)"}
    if roof["frac"] > 1.0:
        roof["note"] = "the peak is a COPY figure (bytes read + written per second); this kernel only reads, and a read-only stream runs above it"
    tr = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr):
        try:
            t = json.load(open(tr))
            roof["traffic"] = t.get(f"{wl}:{dom}")
            roof["traffic_source"] = "dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture (profiles/traffic.json), not measured in this run"
        except Exception:
            pass
    return roof


def free_device_memory(L):
    import gc

    import torch
    gc.collect()
    torch.cuda.synchronize()
    L.bgb_release_cached_memory()
    torch.cuda.empty_cache()


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override rows (per shard for MG)")
    ap.add_argument("--groups", type=int, default=0, help="override key cardinality")
    ap.add_argument("--cpu-sample-rows", type=int, default=50_000_000)
    ap.add_argument("--verify-rows", type=int, default=50_000_000, help="rows of the prefix checked per group against the oracle")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg")
    ap.add_argument("--no-sweep", action="store_true", help="headline workload only")
    ap.add_argument("--no-verify", action="store_true", help="skip the result checks (experiments only: the line then says verified: null)")
    ap.add_argument("--force-path", type=int, default=0)
    ap.add_argument("--exchange", default="auto", choices=["auto", "peer", "nccl"])
    ap.add_argument("--shards", type=int, default=8, help="MG: total number of shards of the job")
    ap.add_argument("--pieces", type=int, default=0, help="MG: pieces every shard is fed as (0 = auto)")
    ap.add_argument("--fold-rows", type=int, default=0, help="MG: pending partial rows that trigger a fold (0 = auto)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch

    from cudf_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus != world and world > 1:
        args.gpus = world
    wl = default_workload(args)
    if wl == "MG":
        run_multi(args, world)
        return
    if world > 1:
        raise SystemExit("single-GPU workloads run with --gpus 1; --gpus N>1 runs MG (configs[4])")
    device = torch.device("cuda", 0)
    torch.cuda.set_device(device)
    L = _lib.lib()
    peak, peak_src = peaks()
    n, g, dtype, desc = WORKLOADS[wl]
    n = args.rows or n
    g = args.groups or g

    # ---- headline
    clocks = ClockSampler(0)
    clocks.start()
    m = measure_single(wl, n, g, 42, device, args.steps, args.warmup, args.force_path, L)
    if not clocks.samples:
        clocks.sample()
    clk = clocks.stop()
    value = n / (m["ms_per_step"] * 1e-3)
    roof = roofline_of(wl, n, g, m, peak, peak_src)
    log(f"{wl}: {m['ms_per_step']:.3f} ms/step  path={m['path']}  kernels={ {k: round(v, 3) for k, v in m['kernels_ms'].items()} }")
    verified, vdetails = None, None
    if not args.no_verify:
        verified, vdetails = verify_workload(wl, m["keys"], m["reqs"], n, g, 42, args.verify_rows, args.force_path)
        log(f"{wl}: verified={verified} {vdetails}")
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(m["keys"], m["reqs"], n, args)
        log(f"{wl}: e2e {e2e['ms_per_step']:.1f} ms/step")
    alg = algorithmic_bytes(wl, n, g)
    del m["keys"], m["reqs"]
    free_device_memory(L)

    # ---- sweep: the other single-GPU workloads, same method, fewer steps
    sweep = None
    if not args.workload and not args.no_sweep:
        sweep = []
        for sw in SWEEP:
            sn, sg, sdtype, sdesc = WORKLOADS[sw]
            sm = measure_single(sw, sn, sg, 42, device, max(3, min(args.steps, 5)), 3, 0, L)
            sroof = roofline_of(sw, sn, sg, sm, peak, peak_src)
            ent = {"workload": sw, "description": sdesc, "rows": sn, "groups": sg, "dtype": sdtype, "ms_per_step": sm["ms_per_step"],
                   "value": sn / (sm["ms_per_step"] * 1e-3), "path": sm["path"], "kernel": sroof["kernel"] if sroof else None,
                   "kernel_ms": sroof["kernel_ms"] if sroof else None, "frac": sroof["frac"] if sroof else None,
                   "whole_call_frac": sroof["whole_call"]["frac"] if sroof else None,
                   "kernels_ms": sroof["kernels_ms"] if sroof else None, "gpu_launches": sm["launches"]}
            if not args.no_verify:
                ok, det = verify_workload(sw, sm["keys"], sm["reqs"], sn, sg, 42, args.verify_rows)
                ent["verified"] = ok
                ent["verify"] = det
            log(f"sweep {sw}: {sm['ms_per_step']:.3f} ms  path={sm['path']}  frac={ent['frac']}  verified={ent.get('verified')}")
            sweep.append(ent)
            del sm
            free_device_memory(L)

    cpu = None
    pandas_b = None
    if not args.no_cpu:
        v, cores, sample, secs = cpu_leg(wl, n, g, 42, args.cpu_sample_rows)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "seconds": secs}
        pandas_b = pandas_leg(wl, n, g, 42, min(args.cpu_sample_rows, 20_000_000))

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": dtype, "data": "synthetic",
        "config": {"workload": f"{wl}: {desc}", "rows_per_gpu": n, "groups": g, "path": m["path"],
                   "l2": "inputs (%.1f GB per step) are larger than the 126 MB L2; no flush needed" % (alg / 1e9),
                   "parallelism": "single"},
        "verified": verified, "verify": vdetails,
        "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": m["launches"], "clocks": clk,
        # host wall-clock of each call (every call ends in a stream sync, so this tracks device time): spread of the steps
        "step_ms": {"median": statistics.median(m["host_ms"]), "min": min(m["host_ms"]), "max": max(m["host_ms"])},
    }
    if sweep is not None:
        line["sweep"] = sweep
    if pandas_b:
        line["pandas_baseline"] = pandas_b
    emit_line(line)


def _pinned_like(col):
    """device Column -> pinned host tensors (data, mask words)"""
    import torch
    d = torch.empty(col.data.shape, dtype=col.data.dtype, pin_memory=True)
    d.copy_(col.data)
    m = None
    if col.mask is not None:
        m = torch.empty(col.mask.shape, dtype=col.mask.dtype, pin_memory=True)
        m.copy_(col.mask)
    return d, m


def run_e2e(keys, reqs, n, args):
    """The same metric through bgb_groupby_host: pinned HOST columns in, host results out; the H2D copy of every input byte
    and the D2H copy of the result are inside the timed region."""
    import torch

    import cudf_b200 as cb
    hold = []

    def host_view(col):
        d, m = _pinned_like(col)
        hold.append((d, m))
        return cb.HostColumnView(d.numpy(), m.numpy().view(np.uint32) if m is not None else None, col.null_count, col.type_id)

    hkeys = [host_view(c) for c in keys.columns()]
    hreqs = [(host_view(r._values), r._aggregations) for r in reqs]
    torch.cuda.synchronize()
    h2d = sum(d.numel() * d.element_size() + (m.numel() * 4 if m is not None else 0) for d, m in hold)
    steps = max(1, min(args.steps, 3))
    for _ in range(2):  # two untimed calls: the library's pinned result buffers are cached once two result sets exist (the previous
        out = cb.groupby_host(hkeys, hreqs, options={"force_path": args.force_path})   # result is still referenced while the next call runs)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        out = cb.groupby_host(hkeys, hreqs, options={"force_path": args.force_path})
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / steps
    G, kout, rout = out
    d2h = sum(a[0].nbytes + (a[1].nbytes if a[1] is not None else 0) for a in kout) + \
        sum(a[0].nbytes + (a[1].nbytes if a[1] is not None else 0) for r in rout for a in r)
    return {"value": n / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": dt * 1e3,
            "steps": steps, "groups_out": int(G), "api": "bgb_groupby_host (C ABI, pinned host columns)"}


# ------------------------------------------------------------------------------------------------ configs[4]: strong scaling
def run_multi(args, world):
    """BASELINE configs[4] at any number of GPUs (see the module docstring): the job is `--shards` shards of `rows` rows."""
    import torch

    import cudf_b200 as cb
    from cudf_b200 import _lib
    from cudf_b200 import distributed as D
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world == 1:
        # one GPU holds all 128 GB of inputs: the library's pool and its block cache are bounded so that torch's allocator (outputs,
        # concatenations) and the pool do not starve each other of the remaining ~60 GB
        pass
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()
    peak, peak_src = peaks()
    _, g, dtype, desc = WORKLOADS["MG"]
    g = args.groups or g
    shard_rows = args.rows or 1_000_000_000
    nshards = args.shards
    if nshards % world != 0:
        raise SystemExit(f"--shards {nshards} must be a multiple of the number of ranks ({world})")
    mine = list(range(rank * nshards // world, (rank + 1) * nshards // world))
    # one GPU holding all 8 shards (128 GB of its 191) has ~60 GB left for scratch, partial tables and folds: every shard is fed as
    # halves (a 5e8-row call: 19 GB of record buffers + 5 GB of table). Smaller pieces cost more than they save -- a quarter
    # shard still meets 9.2e7 of the 1e8 keys, so its call pays the full table and its partial table is as large as a half's
    # (measured r02: quarters 586-744 ms/step, halves 391 ms/step)
    pieces_per_shard = args.pieces or (2 if world == 1 and shard_rows * nshards > 4_000_000_000 else 1)
    shards, total_sum, total_rows = [], 0, 0
    for sidx in mine:
        keys, reqs = make_device_inputs("MG", shard_rows, g, 42 + 2 * sidx, device)  # (seed, seed + 1) per shard: key and value streams never collide
        total_sum += int(reqs[0]._values.data.sum(dtype=torch.int64).item())
        total_rows += shard_rows
        if pieces_per_shard == 1:
            shards.append((keys, reqs))
        else:
            h = shard_rows // pieces_per_shard
            for pc in range(pieces_per_shard):
                lo = pc * h
                sz = h if pc < pieces_per_shard - 1 else shard_rows - lo
                shards.append((cb.Table([c.sliced(lo, sz) for c in keys.columns()]),
                               [cb.GroupByRequest(r._values.sliced(lo, sz), r._aggregations) for r in reqs]))
    torch.cuda.synchronize()
    region = 64 << 20
    if world > 1:
        # receive region: every source sends ~G / world^2 records of 20 bytes to this rank; x2 headroom
        region = max(region, int(2.0 * 20 * g / world) + (8 << 20))

    comm = D.TorchComm() if dist is not None else D.SingleComm()   # ONE communicator: its peer-memory context is created once
    fold_rows = args.fold_rows or (290_000_000 if world == 1 else 450_000_000)          # one GPU holding 128 GB of inputs keeps fewer partial tables pending

    def make_dgb(profile):
        return D.DistributedGroupBy(shards[0][0], options={"profile": profile, "force_path": args.force_path, "fold_rows": fold_rows},
                                    exchange=args.exchange, peer_region_bytes=region, comm=comm)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    dgb = make_dgb(0)
    out = None
    for _ in range(max(args.warmup, 3) if world > 1 else max(min(args.warmup, 3), 1)):
        out = dgb.aggregate_shards(shards)
    barrier()
    launches0 = L.bgb_kernel_launches()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = args.steps if world > 1 else max(1, min(args.steps, 5))
    barrier()
    e0.record()
    for _ in range(steps):
        out = dgb.aggregate_shards(shards)
    e1.record()
    if rank == 0 and not clocks.samples:
        clocks.sample()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = L.bgb_kernel_launches() - launches0
    clk = clocks.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([total_ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / steps
    log(f"rank {rank}: {ms_per_step:.1f} ms/step; torch reserved {torch.cuda.memory_reserved(device) / 1e9:.1f} GB, allocated {torch.cuda.memory_allocated(device) / 1e9:.1f} GB, "
        f"max allocated {torch.cuda.max_memory_allocated(device) / 1e9:.1f} GB, allocator retries {torch.cuda.memory_stats(device).get('num_alloc_retries', 0)}")
    n_total = shard_rows * nshards
    value = n_total / (ms_per_step * 1e-3)

    # ---- verification of the last result: every key once over all ranks, counts and sums add up to the whole job
    okeys, ores = out
    G_local = okeys.num_rows()
    cnt = int(ores[0].columns()[1].data.sum(dtype=torch.int64).item())
    sm = int(ores[0].columns()[0].data.sum(dtype=torch.int64).item())
    uniq = int(torch.unique(okeys.columns()[0].data).numel()) == G_local
    stats = torch.tensor([G_local, cnt, sm, total_rows, total_sum, int(uniq)], device=device, dtype=torch.int64)
    if dist is not None:
        allst = [torch.zeros_like(stats) for _ in range(world)]
        dist.all_gather(allst, stats)
        stats_all = torch.stack(allst).cpu().numpy()
    else:
        stats_all = stats.cpu().numpy()[None, :]
    G_total = int(stats_all[:, 0].sum())
    verified = bool(stats_all[:, 1].sum() == stats_all[:, 3].sum() and stats_all[:, 2].sum() == stats_all[:, 4].sum() and stats_all[:, 5].all()
                    and 0 < G_total <= g)
    # owner partition: the ranks' key sets are disjoint iff the union of all keys has as many elements as the ranks hold together
    if dist is not None:
        gmax = int(stats_all[:, 0].max())
        padded = torch.full((gmax,), -1, dtype=torch.int64, device=device)
        padded[:G_local] = okeys.columns()[0].data
        gathered = [torch.empty(gmax, dtype=torch.int64, device=device) for _ in range(world)] if rank == 0 else None
        dist.gather(padded, gathered, dst=0)
        if rank == 0:
            allk = torch.cat([gathered[r][:int(stats_all[r, 0])] for r in range(world)])
            verified = verified and int(torch.unique(allk).numel()) == G_total
            del allk, gathered
        del padded
    expected_groups = None

    # ---- phases (separate profiled pass) and the exchange's NVLink rate
    pg = make_dgb(1)
    pg.aggregate_shards(shards)
    pg.aggregate_shards(shards)
    prof = {}
    for name, ms in pg.last_profile:
        prof[name] = prof.get(name, 0.0) + ms
    barrier()
    rec_bytes = 20
    out_bytes = None
    nvlink = None
    if world > 1:
        # every rank holds ~all g groups after its local fold and keeps 1/world of them: (world-1)/world * G_fold records leave
        g_fold = min(g, shard_rows * len(mine))
        out_bytes = rec_bytes * g_fold * (world - 1) / world
        ps = prof.get("peer_scatter")
        if ps:
            nvlink = {"bytes_out_per_gpu": out_bytes, "peer_scatter_ms": ps, "GBps_out_per_gpu": out_bytes / (ps * 1e-3) / 1e9,
                      "peak_GBps": 900.0, "measured_peer_copy_GBps": 770.0}
        elif "peer_signal" in prof:
            # fused exchange: the records leave inside the slice kernel of the last local call (peer stores from its flush), so the
            # transfer has no kernel of its own to time; the rate below is the floor "bytes / whole slice kernel"
            sl = prof.get("hc_slices")
            nvlink = {"bytes_out_per_gpu": out_bytes, "fused_into": "hc_slices", "hc_slices_ms": sl,
                      "GBps_out_per_gpu_floor": (out_bytes / (sl * 1e-3) / 1e9) if sl else None, "peak_GBps": 900.0, "measured_peer_copy_GBps": 770.0}
    # dominant local kernel against the HBM roofline (per call of the local path: one piece of one shard)
    local = {k: v for k, v in prof.items() if not k.startswith("dist_") and not k.startswith("peer_")}
    calls = len(shards)
    roof = None
    if local:
        dom = max(local, key=local.get)
        piece_rows = shards[0][0].num_rows()
        alg = algorithmic_bytes("MG", piece_rows, min(g, piece_rows))
        kms = local[dom] / calls
        ach = alg / (kms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "peak_source": peak_src,
                "kernel_ms": kms, "algorithmic_bytes": alg, "traffic": None,
                "whole_call": {"achieved": algorithmic_bytes("MG", n_total / world, g) / (ms_per_step * 1e-3) / 1e9,
                               "frac": algorithmic_bytes("MG", n_total / world, g) / (ms_per_step * 1e-3) / 1e9 / peak},
                "kernels_ms": {k: round(v, 4) for k, v in prof.items()},
                "kernels_note": "sums over the rank's local calls in one step, from a separate profiled pass; dist_* are whole phases"}
        try:
            roof["traffic"] = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(f"MG:{dom}")
            roof["traffic_source"] = "one ncu --set full capture of the same kernel on one 1e9-row shard (profiles/traffic.json), not measured in this run"
        except Exception:
            pass

    # ---- e2e: the same job through DistributedGroupBy with HOST-resident shards: H2D of every shard inside the timed region,
    # D2H of the rank's result columns at the end of the step
    e2e = None
    shards_per_gpu, last_path, last_exchange = len(mine), dgb.last_path, dgb.last_exchange
    if not args.no_e2e:
        # pinned host copies of (at most 4 of) the pieces, then the device-resident inputs are released: the e2e leg stages from the host
        keep = min(len(shards), 4)
        host = []
        for keys, reqs in shards[:keep]:
            kc, vc = keys.columns()[0], reqs[0]._values
            hk = torch.empty(kc.size, dtype=kc.data.dtype, pin_memory=True)
            hv = torch.empty(vc.size, dtype=vc.data.dtype, pin_memory=True)
            hk.copy_(kc.data[kc.offset:kc.offset + kc.size])
            hv.copy_(vc.data[vc.offset:vc.offset + vc.size])
            host.append((hk, hv))
        piece_rows = [sh[0].num_rows() for sh in shards]
        aggs = shards[0][1][0]._aggregations
        torch.cuda.synchronize()
        del shards, keys, reqs, out, okeys, ores, dgb, pg
        free_device_memory(L)
        e2e = run_e2e_multi(args, world, rank, device, dist, host, piece_rows, aggs, n_total, region, comm)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": max(args.warmup, 3) if world > 1 else max(min(args.warmup, 3), 1),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": dtype, "data": "synthetic",
            "config": {"workload": f"MG: {desc}", "rows_total": n_total, "shards": nshards, "rows_per_shard": shard_rows, "groups": g,
                       "shards_per_gpu": shards_per_gpu, "pieces_per_shard": pieces_per_shard, "path": last_path, "exchange": last_exchange,
                       "l2": "inputs (16 GB per shard) are larger than the 126 MB L2; no flush needed",
                       "parallelism": f"shard{world}: local pre-aggregation per shard, fold on the rank, owner-rank exchange ({last_exchange}), merge"},
            "verified": verified,
            "verify": {"groups_total": G_total, "expected_groups": expected_groups, "rows_counted": int(stats_all[:, 1].sum()),
                       "rows_total": int(stats_all[:, 3].sum()), "sums_match": bool(stats_all[:, 2].sum() == stats_all[:, 4].sum()),
                       "keys_unique_per_rank": bool(stats_all[:, 5].all()),
                       "method": "sum of counts == rows of the job, sum of sums == sum of every value column (int64, exact), every rank's keys unique, "
                                 "the union of all ranks' keys (gathered on rank 0) has as many elements as the ranks hold together (owner sets are "
                                 "disjoint); per-group parity of the same code path against the oracle is in tests/"},
            "roofline": roof, "nvlink": nvlink, "cpu_baseline": None, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
        }
        emit_line(line)
    if dist is not None:
        dist.destroy_process_group()


def run_e2e_multi(args, world, rank, device, dist, host, piece_rows, aggs, n_total, region, comm):
    """End to end at N GPUs: every piece of this rank's shards starts in PINNED HOST memory (`host`: the pinned copies, cycled when
    the rank has more pieces than copies -- pinning 128 GB for the N=1 leg would take minutes; the bytes copied and the device
    work per step are those of the whole job either way); a step copies each piece to the device (H2D inside the timed region,
    piece i + 1 while piece i is aggregated), runs the whole distributed groupby (local aggregation, fold, exchange, merge) and
    copies the rank's result columns back to the host."""
    import torch

    import cudf_b200 as cb
    from cudf_b200 import distributed as D
    keep = len(host)
    npieces = len(piece_rows)

    class HostShards:
        """iterable of (keys, requests): piece i + 1 is copied from pinned host memory on a side stream while piece i is
        aggregated (two device buffers; a piece's call has returned before its buffer is reused)"""
        def __init__(self):
            self.bufs = [(torch.empty(max(piece_rows), dtype=torch.int64, device=device), torch.empty(max(piece_rows), dtype=torch.int64, device=device))
                         for _ in range(2)]
            self.copy_stream = torch.cuda.Stream(device)
        def _issue(self, i):
            hk, hv = host[i % keep]
            dk, dv = self.bufs[i % 2]
            nrow = piece_rows[i]
            self.copy_stream.wait_stream(torch.cuda.current_stream(device))
            with torch.cuda.stream(self.copy_stream):
                dk[:nrow].copy_(hk[:nrow], non_blocking=True)
                dv[:nrow].copy_(hv[:nrow], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
            return ev
        def __iter__(self):
            ev = self._issue(0)
            for i in range(npieces):
                nxt = self._issue(i + 1) if i + 1 < npieces else None
                torch.cuda.current_stream(device).wait_event(ev)
                dk, dv = self.bufs[i % 2]
                nrow = piece_rows[i]
                yield (cb.Table([cb.Column(dk[:nrow])]), [cb.GroupByRequest(cb.Column(dv[:nrow]), aggs)])
                ev = nxt

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    hs = HostShards()
    dgb = D.DistributedGroupBy(None, options={"force_path": args.force_path, "fold_rows": args.fold_rows or (290_000_000 if world == 1 else 450_000_000)},
                               exchange=args.exchange, peer_region_bytes=region, comm=comm)

    def step():
        okeys, ores = dgb.aggregate_shards(hs)
        res = [okeys.columns()[0].data.cpu()] + [c.data.cpu() for c in ores[0].columns()]
        return res

    res = step()
    barrier()
    steps = 2 if world == 1 else max(1, min(args.steps, 3))
    t0 = time.perf_counter()
    for _ in range(steps):
        res = step()
    barrier()
    dt = (time.perf_counter() - t0) / steps
    if dist is not None:
        t = torch.tensor([dt], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    h2d = 16 * sum(piece_rows)
    d2h = sum(x.numel() * x.element_size() for x in res)
    return {"value": n_total / dt, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": dt * 1e3,
            "steps": steps, "groups_out_this_rank": int(res[0].numel()),
            "api": "DistributedGroupBy.aggregate_shards over pinned host pieces (per-rank bytes; max over ranks of the wall time, barrier on both sides)"}


if __name__ == "__main__":
    main()
