python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s1_tests.log; tail -5 gpurun_out/s1_tests.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/s1_bench.json 2> gpurun_out/s1_bench.err; tail -c 2500 gpurun_out/s1_bench.json; tail -3 gpurun_out/s1_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/s1_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s1_ncu.log 2>&1
