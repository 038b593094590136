#!/bin/bash
# cardinality sweep on the A9 shape (1e9 rows, int64 key, SUM+COUNT): automatic path vs the generic kernel forced
for g in 20000 30000 100000 300000 1000000 3000000; do
  for fp in 0 1; do
    python bench.py --workload A9 --steps 3 --warmup 2 --groups $g --force-path $fp --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms']; print('G=$g force=$fp', round(d['ms_per_step'],2), d['config']['path'], {a:round(b,2) for a,b in k.items() if b>0.3})"
  done
done
