#!/bin/bash
# usage: tools/bucket_run.sh [tag] -- bucket-path tests + A9 (1e8 groups) and B benches, one line per bench
tag=${1:-x}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_bucket_path.py -x -q > gpurun_out/${tag}_bucket_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/${tag}_bucket_pytest.log
for cfg in "A9 100000000" "B 0"; do
  set -- $cfg; extra=""; [ "$2" != "0" ] && extra="--groups $2"
  timeout 200 python bench.py --workload $1 $extra --steps 3 --warmup 2 --no-cpu --no-e2e > gpurun_out/${tag}_bench_$1.json 2> gpurun_out/${tag}_bench_$1.err; echo "bench $1 rc=$?"
  python -c "
import json,sys
d=json.loads(open('gpurun_out/${tag}_bench_$1.json').read().strip().splitlines()[-1])
print(round(d['ms_per_step'],2), d['config']['path'], d['roofline']['kernels_ms'])"
  tail -2 gpurun_out/${tag}_bench_$1.err
done
