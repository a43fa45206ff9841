#!/bin/bash
# ticket-group sweep at the chain-parallel shapes (1024 x 32 per GPU of the 8-GPU run, 1024 x 256)
for shape in "1024 32" "1024 64" "1024 256"; do
  for G in 0 2 4 8 16 32 64; do
    GPB_TICKET_GROUP=$G python tools/group_sweep.py $shape 2>&1 | tail -1
  done
done
for P in 0 1; do GPB_CHAIN_PAUSE=$P python tools/trace_analyze.py 1024 32 2>&1 | grep -A12 "fused"; done
