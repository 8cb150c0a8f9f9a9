# round-end evidence (v7): tests, bench (both arms), ncu launch list, ncu --set full of the forward and backward kernels
python -m pytest tests -m gpu -x -q 2>&1 | tail -3 > gpurun_out/g_tests.log; cat gpurun_out/g_tests.log
( time timeout 900 python bench.py > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err ) 2>&1 | grep real; tail -2 gpurun_out/g_bench.err
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/g_ref.json 2> gpurun_out/g_ref.err ) 2>&1 | grep real
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g_launches.csv python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/g_ncu.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"enc16_fused_kernel|vq_tc_kernel|quantize_emit_kernel|vq_resolve_kernel|decoder_code_table_kernel" --launch-skip 12 -c 5 -o gpurun_out/g_full -f python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/g_ncufull.log 2>&1; tail -2 gpurun_out/g_ncufull.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"umma_kernel|scatter_add_kernel" --launch-skip 10 -c 7 -o gpurun_out/g_full_bwd -f python tools/train_probe.py > gpurun_out/g_ncufull_bwd.log 2>&1; tail -2 gpurun_out/g_ncufull_bwd.log
python tools/train_probe.py | tail -20 > gpurun_out/g_train_kernels.log; cat gpurun_out/g_train_kernels.log
