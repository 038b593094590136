# cardinality sweep on the config-A shape (1e8 rows, int64 key, SUM+COUNT)
for g in 4 100 1000 10000 14000 20000 50000 100000 1000000 10000000; do
python bench.py --steps 10 --groups $g --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms']; print('G=$g', round(d['ms_per_step'],3), d['config']['path'], d['roofline']['kernel'], round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3))"
done
