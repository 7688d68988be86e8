#!/bin/bash
# cost-ordered launches: parity test, then A/B through tools/order_ab.py
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity2.py -q -m gpu -k "cost_ordered or sample_pool" -x 2>&1 | tail -n 8
timeout 900 python tools/order_ab.py 2>/dev/null | grep "^{" | tee gpurun_out/order_ab.jsonl | cut -c1-260
