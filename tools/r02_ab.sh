#!/bin/bash
# Round-2 GPU session r: tail-gated row-ahead, 10 CTAs/SM; builder tests after the beam.cuh cleanup.
out=gpurun_out; mkdir -p $out
L=$out/r02r_ab.log; : > $L
timeout 600 python -m pytest tests/test_gpu_builder.py -q -m gpu > $out/r02r_pytest_builder.log 2>&1; echo "pytest builder rc=$?" >> $L; tail -2 $out/r02r_pytest_builder.log >> $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
for v in base tra mb10 tra10 base tra; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 2>&1 | tail -1 >> $L; done
for v in base tra; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run ${v}_sat 64 10 2>&1 | tail -1 >> $L; done
for ef in 16 128 256; do for v in base tra; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v $ef 2>&1 | tail -1 >> $L; done; done
for v in base tra; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run ${v}_2500 64 -2500 2>&1 | tail -1 >> $L; done
cat $L
