#!/bin/bash
# Round-2 GPU session aa: hb_knns on pageable caller buffers through page-locked staging with threaded host copies (HB_HOST_STAGING=1).
out=gpurun_out; mkdir -p $out
L=$out/r02aa_host_staging_threads.log; : > $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
for rep in 1 2 3; do for m in 0 1; do HB_HOST_STAGING=$m python tools/ab_e2e.py mode$m 1 2>&1 | tail -1 >> $L; done; done
for m in 0 1; do HB_HOST_STAGING=$m python tools/ab_e2e.py mode$m 10 2>&1 | tail -1 >> $L; done
HB_HOST_STAGING=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -q -m gpu -x > $out/r02aa_pytest.log 2>&1; echo "pytest (HB_HOST_STAGING=1) rc=$?" >> $L; tail -2 $out/r02aa_pytest.log >> $L
cat $L
