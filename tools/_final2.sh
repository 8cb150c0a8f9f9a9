python -m pytest tests -m gpu -x -q 2>&1 | tail -2
( time timeout 900 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/f_bench.err
python -c "
import json;d=json.load(open('gpurun_out/f_bench.json'));print(d['value'],d['ms_per_step'],d['e2e']['value']);print(d['roofline']);print(d['cpu_baseline'])"
