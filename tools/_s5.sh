python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s5_tests.log; tail -4 gpurun_out/s5_tests.log
timeout 120 python tools/enc_probe.py 2>&1 | tail -3
LIPVQ_NVCC_EXTRA=-DENC_PROFILE python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py > gpurun_out/s5_prof.log 2>&1; head -3 gpurun_out/s5_prof.log
