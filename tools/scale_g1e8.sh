#!/bin/bash
# usage: tools/scale_g1e8.sh N -- BASELINE config 5 shape (A9 with 1e8 groups per shard set) at N GPUs, phases printed
export BGB_PEER_TIMEOUT_MS=${BGB_PEER_TIMEOUT_MS:-15000}
n=$1; mkdir -p gpurun_out
out=gpurun_out/scale_A9_g100000000_n$n.json
if [ "$n" = "1" ]; then
  timeout 500 python bench.py --gpus 1 --steps 3 --warmup 3 --workload A9 --groups 100000000 --no-cpu --no-e2e > $out 2> gpurun_out/scale_g1e8_n$n.err
else
  timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps 3 --warmup 3 --workload A9 --groups 100000000 --no-cpu --no-e2e > $out 2> gpurun_out/scale_g1e8_n$n.err
fi
echo "rc=$?"; tail -3 gpurun_out/scale_g1e8_n$n.err | cut -c1-300
python - <<PY
import json
d=json.loads(open("$out").read().strip().splitlines()[-1])
print("n=$n ms/step=%.2f rows/s=%.3e path=%s" % (d["ms_per_step"], d["value"], d["config"]["path"]))
print({k:round(v,2) for k,v in d["roofline"]["kernels_ms"].items() if v>0.2})
PY
