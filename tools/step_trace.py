"""Host-side breakdown of one (distributed) groupby step: run under torchrun with BGB_TRACE=1 to get the library's phase
times on stderr next to the wall time of the whole Python call. Diagnostics only."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import cudf_b200 as cb  # noqa: E402
from cudf_b200 import distributed as D  # noqa: E402


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    device = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(device)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=device)
    wl = sys.argv[1] if len(sys.argv) > 1 else "A"
    exchange = sys.argv[2] if len(sys.argv) > 2 else "auto"
    profile = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    n, g, _ = bench.WORKLOADS[wl]
    if len(sys.argv) > 4:
        g = int(sys.argv[4])
    keys, reqs = bench.make_device_inputs(wl, n, g, 42 + rank, device)
    op = D.DistributedGroupBy(keys, options={"profile": profile}, exchange=exchange) if world > 1 else cb.GroupBy(keys, options={"profile": profile})
    for it in range(8):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()
        t0 = time.perf_counter()
        op.aggregate(reqs)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        if rank == 0:
            print(f"[step {it}] wall {1e6 * (t1 - t0):.0f} us  path={op.last_path}", file=sys.stderr, flush=True)
    # back-to-back steps without barriers (what bench.py times)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 20 if g < 1_000_000 else 4
    for it in range(reps):
        op.aggregate(reqs)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    if rank == 0:
        print(f"[back-to-back] {1e6 * (t1 - t0) / reps:.0f} us/step", file=sys.stderr, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
