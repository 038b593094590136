import numpy as np, zlib, sys, subprocess, os
sys.path.insert(0, '.')
if len(sys.argv) == 1:
    for case in ["int16:0:2", "int16:1:2", "int16:0:0", "int8:0:2", "float32:0:2", "int16nonull:0:2", "int16novmask:0:2"]:
        r = subprocess.run([sys.executable, __file__, case], capture_output=True, text=True, env=dict(os.environ, CUDA_LAUNCH_BLOCKING="1", BGB_DEBUG="1"))
        print("CASE", case, "rc", r.returncode, (r.stdout + r.stderr).strip().splitlines()[-3:])
    sys.exit(0)
from oracle import oracle as orc
from tests import gpu_util
H = orc.HostColumn
kt, inc, fp = sys.argv[1].split(":")
nonull = kt.endswith("nonull"); novmask = kt.endswith("novmask")
kt = kt.replace("nonull", "").replace("novmask", "")
rng = np.random.default_rng(1)
n = 400_000
keys = [H(rng.integers(0, 3000 if kt != "int8" else 100, n).astype(kt), None if nonull else rng.random(n) > 0.03)]
vals = [H(rng.integers(-1000, 1001, n).astype("int64"), None if novmask else rng.random(n) > 0.1)]
G, path = gpu_util.compare_with_oracle(keys, vals, [(0, "sum", 1), (0, "count_valid", 1)], include_null_keys=bool(int(inc)), options={"force_path": int(fp)}, name="dbg")
print("ok", G, path)
