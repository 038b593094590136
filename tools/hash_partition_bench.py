"""Times hash_partition (bgb_hash_partition) on device-resident columns: python tools/hash_partition_bench.py [rows]"""
import os
import sys

import numpy as np  # noqa: F401
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import partitioning as part

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 100_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)
k64 = torch.randint(0, 1 << 40, (n,), dtype=torch.int64, device=dev, generator=g)
v64 = torch.randint(0, 1000, (n,), dtype=torch.int64, device=dev, generator=g)
k32 = k64.to(torch.int32)
cases = {
    "int64 key + int64 value": ([cb.Column(k64), cb.Column(v64)], [0]),
    "int32 key + int64 value": ([cb.Column(k32), cb.Column(v64)], [0]),
    "2 keys (int64,int32) + int64 value": ([cb.Column(k64), cb.Column(k32), cb.Column(v64)], [0, 1]),
}
for name, (cols, idx) in cases.items():
    bytes_moved = sum(2 * c.data.element_size() for c in cols) + sum(cols[i].data.element_size() for i in idx)
    for nparts in (8, 64, 512, 4096):
        t = cb.Table(cols)
        for _ in range(2):
            part.hash_partition(t, idx, nparts)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        e0.record()
        for _ in range(reps):
            part.hash_partition(t, idx, nparts)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"{name:40s} parts={nparts:5d}  {ms:8.3f} ms  {n / ms / 1e6:8.2f} G rows/s  {bytes_moved * n / ms / 1e6:8.1f} GB/s algorithmic", flush=True)
