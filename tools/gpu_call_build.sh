#!/bin/bash
python -m pytest tests -x -q -m gpu 2>&1 | tail -4
python tools/c3_breakdown.py 2>&1 | head -8
python bench.py --steps 50 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().split('\n')[-1]); print('value',d['value'],'e2e',d['e2e']['value'],'ms/step',d['ms_per_step'],'roofline',d['roofline']['achieved'])"
