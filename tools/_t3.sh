python -m pytest tests -m gpu -x -q -k "trainer" 2>&1 | grep -E "Error|error|assert |passed|failed|^E " | cut -c1-300 | head -20
