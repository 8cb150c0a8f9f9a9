for v in 2 3 4; do LIPVQ_E2E_DEPTH=$v python bench.py --steps 30 --warmup 5 --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('depth=$v', d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step'])"; done
