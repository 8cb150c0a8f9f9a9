set -x
C="655360 208 1024 collapsed 655360 208 1024 uniform 1048576 64 1024 uniform 262144 256 16384 uniform"
LIPVQ_TC_ELEM=1 timeout 200 python tools/vq_probe.py $C 2>&1 | tail -4
LIPVQ_TC_ELEM=0 timeout 200 python tools/vq_probe.py $C 2>&1 | tail -4
timeout 100 python tools/vq_probe.py 655360 208 1024 collapsed 1048576 64 1024 uniform > gpurun_out/r6_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r6_launches.csv python tools/vq_probe.py 655360 208 1024 collapsed 1048576 64 1024 uniform > gpurun_out/r6_ncu.log 2>&1
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r6_bench.json 2> gpurun_out/r6_bench.err; tail -c 1500 gpurun_out/r6_bench.json; tail -3 gpurun_out/r6_bench.err
