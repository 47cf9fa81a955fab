#!/bin/bash
# Round-2 GPU session u: accept loop with contains() folded into the eviction reduction (HB_FUSE_ACCEPT 0/1).
out=gpurun_out; mkdir -p $out
L=$out/r02u_fuse_accept.log; : > $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
for rep in 1 2; do
for v in base fuse; do
  HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 1 2>&1 | tail -1 >> $L
  HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 10 2>&1 | tail -1 >> $L
  HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 128 1 2>&1 | tail -1 >> $L
done; done
timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu > $out/r02u_pytest.log 2>&1; echo "pytest (default lib = fuse) rc=$?" >> $L; tail -2 $out/r02u_pytest.log >> $L
cat $L
