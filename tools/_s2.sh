LIPVQ_NVCC_EXTRA=-DENC_PROFILE python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py > gpurun_out/s2_prof.log 2>&1; cat gpurun_out/s2_prof.log
