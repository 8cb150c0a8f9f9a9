python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s3_tests.log; tail -8 gpurun_out/s3_tests.log
timeout 120 python tools/enc_probe.py 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/s3_bench.json 2> gpurun_out/s3_bench.err; tail -c 2600 gpurun_out/s3_bench.json; tail -3 gpurun_out/s3_bench.err
LIPVQ_NVCC_EXTRA=-DENC_PROFILE python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py > gpurun_out/s3_prof.log 2>&1; cat gpurun_out/s3_prof.log
