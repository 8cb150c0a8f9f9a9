python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/s7_tests.log; tail -3 gpurun_out/s7_tests.log
( time timeout 900 python bench.py > gpurun_out/s7_bench.json 2> gpurun_out/s7_bench.err ) 2>&1 | grep real; tail -c 4500 gpurun_out/s7_bench.json; tail -3 gpurun_out/s7_bench.err
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/s7_ref.json 2> gpurun_out/s7_ref.err ) 2>&1 | grep real; cat gpurun_out/s7_ref.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/s7_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s7_ncu.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"enc_fused_kernel|vq_tc_kernel|quantize_emit_kernel|vq_resolve_kernel|dec_kernel" --launch-skip 12 -c 5 -o gpurun_out/s7_full -f python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s7_ncufull.log 2>&1; tail -2 gpurun_out/s7_ncufull.log
