python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/s8_tests.log; tail -3 gpurun_out/s8_tests.log
timeout 900 python bench.py --no-cpu > gpurun_out/s8_bench.json 2> gpurun_out/s8_bench.err; python -c "
import json;d=json.load(open('gpurun_out/s8_bench.json'));print(d['value'],d['ms_per_step'],d['e2e'])"; tail -3 gpurun_out/s8_bench.err
