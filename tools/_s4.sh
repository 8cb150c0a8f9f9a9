ENC_PROBE_ONCE=1 timeout 120 python tools/enc_probe.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"enc_fused_kernel|vq_tc_kernel" -c 2 -o gpurun_out/s4_enc -f env ENC_PROBE_ONCE=1 python tools/enc_probe.py > gpurun_out/s4_ncu.log 2>&1
tail -3 gpurun_out/s4_ncu.log; ls -la gpurun_out/
