#!/bin/bash
rm -f gpurun_out/c14.log
timeout 300 python tools/trace_analyze.py 1024 256 >> gpurun_out/c14.log 2>&1
timeout 300 python tools/trace_analyze.py 4096 8 >> gpurun_out/c14.log 2>&1
