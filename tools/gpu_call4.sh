#!/bin/bash
timeout 300 python tools/diag_2cta.py 1024 > gpurun_out/c4_diag.log 2>&1
timeout 300 python tools/diag_2cta.py 4096 >> gpurun_out/c4_diag.log 2>&1
