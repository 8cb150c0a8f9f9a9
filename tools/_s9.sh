for v in "-DENC_EXP_NOFENCE" "-DENC_EXP_NOSIG" "-DENC_EXP_NOFENCE -DENC_EXP_NOSIG -DENC_EXP_SKIP=3"; do
LIPVQ_NVCC_EXTRA="-DENC_PROFILE $v" python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
echo "exp=$v"; ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py 2>&1 | head -2
done
