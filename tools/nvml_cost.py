import sys, time, threading
sys.path.insert(0, '.')
import torch, bench, cudf_b200 as cb, pynvml as nv
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0)
for name, f in [("clock", lambda: nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), ("maxclock", lambda: nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)),
                ("reasons", lambda: nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))]:
    f(); t0 = time.perf_counter()
    for _ in range(10): f()
    print(name, "ms/call idle", (time.perf_counter() - t0) / 10 * 1e3)
dev = torch.device("cuda", 0)
keys, reqs = bench.make_device_inputs("A", 100_000_000, 10_000, 42, dev)
gb = cb.GroupBy(keys, options={"profile": 0})
for _ in range(5): gb.aggregate(reqs)
def run(n):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): gb.aggregate(reqs)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print("no sampler ms/step", run(200))
stop = threading.Event(); lat = []
def bg(period, what):
    while not stop.wait(period):
        t0 = time.perf_counter(); what(); lat.append((time.perf_counter() - t0) * 1e3)
for label, what in [("clock only", lambda: nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), ("reasons only", lambda: nv.nvmlDeviceGetCurrentClocksThrottleReasons(h))]:
    for period in (0.05, 0.01):
        stop.clear(); lat.clear()
        t = threading.Thread(target=bg, args=(period, what)); t.start()
        r = run(400); stop.set(); t.join()
        print(label, "period", period, "ms/step", r, "nvml ms/call under load", sum(lat) / max(len(lat), 1), len(lat))
