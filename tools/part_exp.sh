timeout 600 python -m pytest tests/test_gpu_partitioned_path.py tests/test_gpu_distributed.py -x -q 2>&1 | tail -3
python bench.py --steps 3 --workload A9 --groups 100000000 --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms']; print(round(d['ms_per_step'],2), k)"
python bench.py --steps 3 --workload B --no-cpu --no-e2e 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); k=d['roofline']['kernels_ms']; print('B', round(d['ms_per_step'],2), k)"
