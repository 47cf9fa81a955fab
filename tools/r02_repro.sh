#!/bin/bash
out=gpurun_out; mkdir -p $out; L=$out/r02k_repro.log; : > $L
for args in "0 0" "87500000 0" "87500000 1"; do
  echo "== repro_100m $args" >> $L
  CUDA_LAUNCH_BLOCKING=1 timeout 600 python tools/repro_100m.py $args >> $L 2>&1; echo "rc=$?" >> $L
done
grep -v "^frame\|^$" $L | tail -60
