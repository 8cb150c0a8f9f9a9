for v in 1 2 3; do
LIPVQ_NVCC_EXTRA="-DENC_PROFILE -DENC_EXP_SKIP=$v" python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
echo "skip=$v"; ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py 2>&1 | head -2
done
