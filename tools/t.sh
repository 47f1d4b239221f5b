timeout 300 python -m pytest tests/test_abi.py -m gpu -q 2>&1 | tail -30
