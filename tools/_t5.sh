python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/s18_bench2.json 2> gpurun_out/s18_bench2.err; python -c "
import json;d=json.load(open('gpurun_out/s18_bench2.json'));print(d['n_gpus'],d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step']); print(d['train_step'])"; tail -3 gpurun_out/s18_bench2.err
