#!/bin/bash
python -m pytest tests/test_gpu_parity.py tests/test_cpp_mirror.py -x -q -m gpu -k "mll_gradient or hmc" 2>&1 | tail -5
python tools/grad_timing.py || exit 1
