#!/bin/bash
# usage: tools/scale_run.sh "<N list>" -- runs bench.py at each N on one box and leaves one JSON line per run in gpurun_out/
export BGB_PEER_TIMEOUT_MS=${BGB_PEER_TIMEOUT_MS:-15000}
mkdir -p gpurun_out
for n in $1; do
  for cfg in "A 50 0" "A9 10 0" "A9 3 100000000"; do
    set -- $cfg; wl=$1; steps=$2; groups=$3
    extra=""; tag=$wl
    if [ "$groups" != "0" ]; then extra="--groups $groups"; tag="${wl}_g$groups"; fi
    out=gpurun_out/scale_${tag}_n$n.json
    if [ "$n" = "1" ]; then
      timeout 400 python bench.py --gpus 1 --steps $steps --warmup 3 --workload $wl $extra --no-cpu --no-e2e > $out 2> gpurun_out/scale_${tag}_n$n.err
    else
      timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --steps $steps --warmup 3 --workload $wl $extra --no-cpu --no-e2e > $out 2> gpurun_out/scale_${tag}_n$n.err
    fi
    echo "n=$n $tag rc=$? $(python - <<PY
import json
try:
    d=json.loads(open("$out").read().strip().splitlines()[-1])
    print("ms/step=%.3f rows/s=%.3e path=%s step_ms=%s" % (d["ms_per_step"], d["value"], d["config"]["path"], d.get("step_ms")))
except Exception as e:
    print("no json:", e)
PY
)"
    tail -2 gpurun_out/scale_${tag}_n$n.err | cut -c1-300
  done
done
