python -m pytest tests -m gpu -x -q 2>&1 | tail -8
python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/s16_bench.json 2> gpurun_out/s16_bench.err; tail -3 gpurun_out/s16_bench.err
python -c "
import json;d=json.load(open('gpurun_out/s16_bench.json'));print(d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step']);print(d['kernels_us_per_step']);print(d['train_step']); print(d['config2_vq_alone']['ms'], d['config2_vq_alone']['kernels_us'])"
