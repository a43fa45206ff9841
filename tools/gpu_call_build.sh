#!/bin/bash
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "loglik_parity or schedulers or full_size_c2 or full_size_c3 or mh_ or predict" 2>&1 | tail -3
python tools/c3_breakdown.py 2>&1 | head -8
python tools/trace_analyze.py 4096 8 2>&1 | grep -A12 "offdiag" | grep "dep ok\|producer"
