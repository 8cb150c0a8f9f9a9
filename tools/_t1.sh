python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 120 python tools/enc_probe.py 2>&1 | tail -2
LIPVQ_ENC=tf32 timeout 120 python tools/enc_probe.py 2>&1 | tail -2
