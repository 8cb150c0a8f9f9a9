python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for v in 1 0; do LIPVQ_SIDE_STREAM=$v python bench.py --steps 30 --warmup 5 --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('side=$v', d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step']);print(d['kernels_us_per_step'])"; done
