"""Float MEAN / SUM+MIN+MAX over a dense integer key range of 1e3-4e3 groups: the tile-sorting kernel (automatic choice)
against the generic HBM-table kernel (force_path=1), plus BASELINE configs[0]'s shape (two int64 keys in [0, 32), two float64
MEANs) at the given row count. python tools/sorted_bench.py [rows]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import _lib, aggregation as A  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
dev = torch.device("cuda:0")
L = _lib.lib()
s = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def gen(seed, mod, add=0):
    t = torch.empty(n, dtype=torch.int64, device=dev)
    _lib.check(L.bgb_gen_uniform_int(C.c_void_p(t.data_ptr()), cb.TypeId.INT64, n, seed, 0, mod, add, s))
    return cb.Column(t)


def fcol(seed):
    g = torch.Generator(device=dev); g.manual_seed(seed)
    return cb.Column(torch.randn(n, dtype=torch.float64, device=dev, generator=g))


def run(name, keys, reqs):
    for fp in (0, 1):
        gb = cb.GroupBy(keys, options={"force_path": fp, "profile": 1})
        for _ in range(2):
            gb.aggregate(reqs)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            kt, _ = gb.aggregate(reqs)
        e1.record(); torch.cuda.synchronize()
        print(f"{name:40s} force={fp} {e0.elapsed_time(e1) / 5:8.3f} ms/step  groups={kt.num_rows()}  path={gb.last_path}  "
              + " ".join(f"{k}={v:.3f}" for k, v in gb.last_profile if v > 0.02), flush=True)


c, d = fcol(11), fcol(12)
for g in (300, 1000, 1900):
    keys = cb.Table([gen(1, g, -g // 2)])
    run(f"1 int64 key, {g} groups, 2 f64 MEAN", keys, [cb.GroupByRequest(c, [A.mean()]), cb.GroupByRequest(d, [A.mean()])])
    run(f"1 int64 key, {g} groups, f64 SUM+MIN+MAX", keys, [cb.GroupByRequest(c, [A.sum(), A.min(), A.max()])])
keys = cb.Table([gen(1, 32), gen(2, 32)])
run("C1: 2 int64 keys 32x32, 2 f64 MEAN", keys, [cb.GroupByRequest(c, [A.mean()]), cb.GroupByRequest(d, [A.mean()])])
# sparse keys: ids come from the CTA's hash table ("sorted-hashed")
for g in (300, 600):
    t = torch.empty(n, dtype=torch.int64, device=dev)
    _lib.check(L.bgb_gen_uniform_int(C.c_void_p(t.data_ptr()), cb.TypeId.INT64, n, 5, 0, g, 0, s))
    t.mul_(0x9E3779B97F4A7C15 - (1 << 64)).bitwise_xor_(t >> 29)   # spread the ids over the whole int64 range
    keys = cb.Table([cb.Column(t)])
    run(f"1 sparse int64 key, {g} groups, 2 f64 MEAN", keys, [cb.GroupByRequest(c, [A.mean()]), cb.GroupByRequest(d, [A.mean()])])
