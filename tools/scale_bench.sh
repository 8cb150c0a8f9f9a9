N=$1; nvidia-smi topo -m 2>/dev/null | head -14; lscpu | grep -E "NUMA|Socket|^CPU\(s\)" 
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/scale_bench_${N}gpu.json 2> gpurun_out/scale_bench_${N}gpu.err
python -c "
import json;d=json.load(open('gpurun_out/scale_bench_${N}gpu.json'));print(d['n_gpus'],d['value'],d['ms_per_step'],d['e2e']['value'],d['e2e']['ms_per_step']); print(d['train_step'])"; tail -2 gpurun_out/scale_bench_${N}gpu.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps 2 --warmup 1 | cut -c1-160
