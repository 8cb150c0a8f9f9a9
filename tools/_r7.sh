C="655360 208 1024 collapsed"
timeout 100 python tools/vq_probe.py $C > gpurun_out/r7_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"vq_rescore|vq_tc_kernel|vq_convert" -s 6 -c 3 -o gpurun_out/r7_prof -f python tools/vq_probe.py $C > gpurun_out/r7_ncu.log 2>&1
tail -3 gpurun_out/r7_ncu.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/r7_bench.json 2> gpurun_out/r7_bench.err; tail -c 1800 gpurun_out/r7_bench.json; tail -3 gpurun_out/r7_bench.err
