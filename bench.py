#!/usr/bin/env python
"""bench.py -- hash groupby-aggregate throughput on B200 (rows/s + fraction of the HBM roofline).

Contract (see the task statement): `python bench.py --gpus N --steps K --warmup W` prints ONE JSON line on rank 0.
A "step" is one full groupby call (build + emit: hash, aggregate, compact, gather/finalise) over one batch of synthetic
columns that are already resident in HBM; `e2e` is the same call through the C ABI's host-buffer entry point
(bgb_groupby_host) with pinned HOST columns, H2D/D2H inside the timed region.

Workloads (BASELINE.json configs / SURVEY.md 8d):
  A      int64 key x%1e4, int64 value, SUM+COUNT, N=1e8            (default at N=1: configs[1])
  A9     as A with N=1e9 and --groups 10000 | 100000000             (north_star target shapes)
  B      int32 key x%1e8, float64 value 10% null, SUM+MIN+MAX, N=1e9
  Q1     2 int8 keys (4 groups), 2 float64 values, MEAN+VAR+COUNT each, N=6e8
At N>1 every rank aggregates its own shard of the same per-GPU workload (weak scaling), the per-rank partial groups are
hash-partitioned by key and exchanged with one NCCL all-to-all, and merged by a second groupby on the owner rank.

`--impl reference` times the CPU restatement of the reference algorithm (oracle/, all host threads) on a bounded sample
of the same workload: libcudf itself cannot be built offline, so the oracle port is the reference arm.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
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
    # name: (rows, groups, description)
    "A": (100_000_000, 10_000, "single int64 key SUM/COUNT, 100M rows, 1e4 groups (BASELINE.json configs[1])"),
    "A9": (1_000_000_000, 10_000, "single int64 key SUM/COUNT, 1e9 rows (north_star target shape)"),
    "B": (1_000_000_000, 100_000_000, "int32 key SUM/MIN/MAX float64 with 10% nulls, 1e9 rows, 1e8 groups (configs[2])"),
    "Q1": (600_000_000, 4, "TPC-H Q1-style two int8 keys MEAN/VAR/COUNT over 6e8 rows (configs[3])"),
    "C1": (10_000_000, 1024, "df.groupby(['A','B']).mean(): two int64 keys in [0,32), two float64 values, 10M rows (configs[0] shape on the GPU)"),
}


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
    if workload in ("A", "A9"):
        return 16.0 * n + 20.0 * g
    if workload == "B":
        return 12.125 * n + (4 + 24 + 3 / 8) * g
    if workload == "Q1":
        return 18.0 * n
    if workload == "C1":
        return 32.0 * n + 32.0 * g
    raise ValueError(workload)


def make_device_inputs(workload: str, n: int, g: int, seed: int, device):
    """Columns generated ON the device by the library's counter-based SplitMix64 generators (bit-identical to the oracle's)."""
    import ctypes as C

    import torch

    import cudf_b200 as cb
    from cudf_b200 import _lib, aggregation as A
    L = _lib.lib()
    s = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)

    def gen_int(dtype, tid, sd, mod, add=0):
        t = torch.empty(n, dtype=dtype, device=device)
        _lib.check(L.bgb_gen_uniform_int(C.c_void_p(t.data_ptr()), tid, n, sd, 0, mod, add, s))
        return t

    def gen_real(sd, mod, add=0.0):
        t = torch.empty(n, dtype=torch.float64, device=device)
        _lib.check(L.bgb_gen_uniform_real(C.c_void_p(t.data_ptr()), cb.TypeId.FLOAT64, n, sd, 0, mod, add, s))
        return t

    if workload in ("A", "A9"):
        k = cb.Column(gen_int(torch.int64, cb.TypeId.INT64, seed, g))
        v = cb.Column(gen_int(torch.int64, cb.TypeId.INT64, seed + 1, 2001, -1000))
        return cb.Table([k]), [cb.GroupByRequest(v, [A.sum(), A.count()])]
    if workload == "B":
        k = cb.Column(gen_int(torch.int32, cb.TypeId.INT32, seed, g))
        vd = gen_real(seed + 1, 1_000_000)
        mask = torch.empty((n + 31) // 32, dtype=torch.int32, device=device)
        nc = C.c_int32(0)
        _lib.check(L.bgb_gen_null_mask(C.c_void_p(mask.data_ptr()), n, seed + 2, 0, 10, C.byref(nc), s))
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
        _lib.check(L.bgb_gen_categorical_i8(C.c_void_p(rf.data_ptr()), n, seed, 0, cum, bytes([0, 1, 1, 2]), 4, s))
        _lib.check(L.bgb_gen_categorical_i8(C.c_void_p(ls.data_ptr()), n, seed, 0, cum, bytes([0, 0, 1, 0]), 4, s))
        q = cb.Column(gen_real(seed + 1, 50, 1.0))
        p = cb.Column(gen_real(seed + 2, 104_100, 900.0))
        aggs = lambda: [A.mean(), A.variance(1), A.count()]  # noqa: E731
        return cb.Table([cb.Column(rf), cb.Column(ls)]), [cb.GroupByRequest(q, aggs()), cb.GroupByRequest(p, aggs())]
    raise ValueError(workload)


def make_host_inputs_numpy(workload: str, n: int, g: int, seed: int):
    """The same columns on the host (numpy), for the CPU legs. Identical values to make_device_inputs."""
    from oracle import oracle as orc
    rows = np.arange(n, dtype=np.uint64)
    H = orc.HostColumn
    if workload in ("A", "A9"):
        k = (orc.splitmix64(seed, rows) % np.uint64(g)).astype(np.int64)
        v = (orc.splitmix64(seed + 1, rows) % np.uint64(2001)).astype(np.int64) - 1000
        return [H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)]
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


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    wl = args.workload or "A"
    n, g, desc = WORKLOADS[wl]
    n = args.rows or n
    g = args.groups or g
    sample = min(n, args.cpu_sample_rows)
    keys, vals, aggs = make_host_inputs_numpy(wl, sample, g, 42)
    from oracle import oracle as orc
    cores = os.cpu_count() or 1
    for _ in range(args.warmup):
        orc.groupby(keys, vals, aggs, threads=cores, sort=False)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        orc.groupby(keys, vals, aggs, threads=cores, sort=False)
    dt = (time.perf_counter() - t0) / args.steps
    value = sample / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64" if wl in ("A", "A9") else "f64", "data": "synthetic",
        "config": {"workload": f"{wl}: {desc}", "rows": n, "groups": g, "sample_rows_per_step": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"each step = first {sample} rows of the workload; libcudf cannot be built offline, "
                                   "so the reference arm is the CPU restatement in oracle/ (OpenMP, all host cores)"},
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


# ------------------------------------------------------------------------------------------------ our arm
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


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, choices=list(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0, help="override rows per GPU")
    ap.add_argument("--groups", type=int, default=0, help="override key cardinality")
    ap.add_argument("--cpu-sample-rows", type=int, default=50_000_000)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg")
    ap.add_argument("--force-path", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        run_reference_arm(args)
        return

    import torch

    import cudf_b200 as cb
    from cudf_b200 import _lib
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.gpus != world and world > 1:
        args.gpus = world
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=device)
    L = _lib.lib()

    wl = args.workload or "A"
    n, g, desc = WORKLOADS[wl]
    n = args.rows or n
    g = args.groups or g
    seed = 42 + rank
    keys, reqs = make_device_inputs(wl, n, g, seed, device)
    torch.cuda.synchronize()
    opts = {"profile": 1, "force_path": args.force_path}
    gb = cb.GroupBy(keys, options=opts)
    dgb = None
    if world > 1:
        from cudf_b200 import distributed as D
        dgb = D.DistributedGroupBy(keys, options=opts)

    def step():
        if dgb is not None:
            return dgb.aggregate(reqs)
        return gb.aggregate(reqs)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    launches0 = L.bgb_kernel_launches()
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof_acc = {}
    barrier()
    e0.record()
    host_ms = []
    for it in range(args.steps):
        th = time.perf_counter()
        step()
        host_ms.append((time.perf_counter() - th) * 1e3)
        src = dgb.last_profile if dgb is not None else gb.last_profile
        for name, ms in src:
            prof_acc.setdefault(name, []).append(ms)
    e1.record()
    if rank == 0 and not clocks.samples:
        # region shorter than the sampler's 20 ms period: one sample while the GPU is still finishing the last step. Taken
        # after e1 is recorded because an NVML query from the launching thread costs ~2 ms (it would be billed to the steps).
        clocks.sample()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = L.bgb_kernel_launches() - launches0
    clk = clocks.stop() if rank == 0 else None
    if dist is not None:
        t = torch.tensor([total_ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * n / (ms_per_step * 1e-3)
    path = (dgb.last_path if dgb is not None else gb.last_path)

    # ---- roofline of the dominant kernel (CUDA events on the call's stream, recorded inside the timed steps)
    peak, peak_src = peaks()
    kern_ms = {k: sum(v) / len(v) for k, v in prof_acc.items()}
    local_ms = {k: v for k, v in kern_ms.items() if not k.startswith("dist_")}  # dist_* are whole phases, not kernels
    dom = max(local_ms, key=local_ms.get) if local_ms else None
    alg = algorithmic_bytes(wl, n, g)
    roof = None
    if dom:
        ach = alg / (kern_ms[dom] * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "peak_source": peak_src, "kernel_ms": kern_ms[dom], "algorithmic_bytes": alg, "traffic": None,
                "frac_of_nominal_8000": ach / 8000.0,
                "whole_call": {"achieved": alg / (ms_per_step * 1e-3) / 1e9, "frac": alg / (ms_per_step * 1e-3) / 1e9 / peak},
                "kernels_ms": {k: round(v, 4) for k, v in kern_ms.items()}}
        tr = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr):
            try:
                roof["traffic"] = json.load(open(tr)).get(f"{wl}:{dom}")
            except Exception:
                pass

    # ---- end to end through the host-buffer C ABI call (pinned host columns; H2D + D2H inside the timed region)
    e2e = None
    if not args.no_e2e and world == 1:
        e2e = run_e2e(wl, keys, reqs, n, args, device)
    elif not args.no_e2e:
        e2e = run_e2e_distributed(wl, keys, reqs, n, args, device, dist, world)

    cpu = None
    pandas_b = None
    if rank == 0 and world == 1 and not args.no_cpu:
        v, cores, sample, secs = cpu_leg(wl, n, g, 42, args.cpu_sample_rows)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, "seconds": secs}
        pandas_b = pandas_leg(wl, n, g, 42, min(args.cpu_sample_rows, 20_000_000))

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int64" if wl in ("A", "A9") else "f64", "data": "synthetic",
            "config": {"workload": f"{wl}: {desc}", "rows_per_gpu": n, "groups": g, "path": path,
                       "l2": "inputs (%.1f GB per step) are larger than the 126 MB L2; no flush needed" % (alg / 1e9),
                       "parallelism": f"shard{world}" if world > 1 else "single"},
            "roofline": roof, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk,
            # host wall-clock of each call (every call ends in a stream sync, so this tracks device time): spread of the steps
            "step_ms": {"median": statistics.median(host_ms), "min": min(host_ms), "max": max(host_ms)},
        }
        if pandas_b:
            line["pandas_baseline"] = pandas_b
        emit_line(line)
    if dist is not None:
        dist.destroy_process_group()


def _pinned_like(col):
    """device Column -> pinned host numpy views (data, mask words)"""
    import torch
    d = torch.empty(col.data.shape, dtype=col.data.dtype, pin_memory=True)
    d.copy_(col.data)
    m = None
    if col.mask is not None:
        m = torch.empty(col.mask.shape, dtype=col.mask.dtype, pin_memory=True)
        m.copy_(col.mask)
    return d, m


def run_e2e(wl, keys, reqs, n, args, device):
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
    steps = max(1, min(args.steps, 5))
    out = None
    for _ in range(1):
        out = cb.groupby_host(hkeys, hreqs, options={"force_path": args.force_path})
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


def run_e2e_distributed(wl, keys, reqs, n, args, device, dist, world):
    """N>1: every rank runs the host-buffer call on its shard, then the distributed exchange+merge of the partials."""
    import torch
    e = run_e2e(wl, keys, reqs, n, args, device)
    t = torch.tensor([e["ms_per_step"]], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    e.update({"value": world * n / (ms * 1e-3), "ms_per_step": ms, "note": "per-rank host-buffer call, max over ranks"})
    return e


if __name__ == "__main__":
    main()
