timeout 250 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
WC_HOST_REDUCE=1 timeout 250 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
WC_TIMING=1 timeout 150 python tools/e2e_probe.py 2>&1 | tail -2
