for X in "" "-DENC_CH_=64" "-DENC_NOSTREAM" "-DENC_CH_=64 -DENC_NOSTREAM" "-DENC_CH_=32"; do
  LIPVQ_NVCC_EXTRA="$X" python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)" && LIPVQ_NVCC_EXTRA="$X" timeout 120 python tools/enc_probe.py
done
python -c "from lipvq_vae_b200 import _native as N; N.build_native(force=True)"
