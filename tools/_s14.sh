python -m pytest tests -m gpu -x -q 2>&1 | tail -6
python tools/train_probe.py | tail -1
LIPVQ_GEMM=simt python tools/train_probe.py | tail -1
