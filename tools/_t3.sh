python -m pytest tests -m gpu -x -q -k fused_encoder 2>&1 | grep -E "AssertionError|assert |passed|failed" | cut -c1-300 | head
