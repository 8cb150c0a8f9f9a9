ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"enc16_fused_kernel" -c 1 -o gpurun_out/s17_enc16 -f env ENC_PROBE_ONCE=1 python tools/enc_probe.py > gpurun_out/s17_ncu.log 2>&1
tail -2 gpurun_out/s17_ncu.log
