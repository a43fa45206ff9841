#!/bin/bash
# A/B timing of the factorisation kernel over the shapes that matter (chol_shape.py: median of 6)
mkdir -p gpurun_out
out=gpurun_out/${1:-ab}.log
: > $out
for shape in "4096 1" "2048 1" "1024 1" "300 2" "1024 32" "1024 64" "1024 256" "4096 2" "4096 8" "8192 8"; do
  timeout 120 python tools/chol_shape.py $shape 6 >> $out 2>&1
done
cat $out
