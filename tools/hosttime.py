import sys, time
sys.path.insert(0, '.')
import torch, bench, cudf_b200 as cb
dev = torch.device("cuda", 0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
keys, reqs = bench.make_device_inputs("A", N, 10_000, 42, dev)
for prof in (0, 1):
    gb = cb.GroupBy(keys, options={"profile": prof})
    for _ in range(3): gb.aggregate(reqs)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(20): gb.aggregate(reqs)
    torch.cuda.synchronize()
    print("profile", prof, "wall ms/step", (time.perf_counter() - t0) / 20 * 1e3, gb.last_path)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): gb.aggregate(reqs)
    e1.record(); torch.cuda.synchronize()
    print("  event ms/step", e0.elapsed_time(e1) / 5, gb.last_profile)
import cProfile, pstats
gb = cb.GroupBy(keys, options={"profile": 0})
pr = cProfile.Profile(); pr.enable()
for _ in range(20): gb.aggregate(reqs)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(12)
