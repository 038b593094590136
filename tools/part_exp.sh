for r in 1 2 4; do for g in 50000; do
BGB_GENERIC_ROWS=$r python bench.py --steps 10 --groups $g --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms']; print('R=$r G=$g', round(d['ms_per_step'],3), d['config']['path'], d['roofline']['kernel'], round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3))"
done; done
