#!/bin/bash
# Round-2 GPU session s: ground-truth warm-start pass (HB_GT_WARM 0 off / 1 k > 16 / 2 always): parity tests, then timings.
out=gpurun_out; mkdir -p $out
L=$out/r02s_gt_warm.log; : > $L
HB_GT_WARM=2 timeout 900 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "ground or truth or gt or tied or smoke or cross" > $out/r02s_pytest_gt_warm2.log 2>&1; echo "pytest (HB_GT_WARM=2) rc=$?" >> $L; tail -2 $out/r02s_pytest_gt_warm2.log >> $L
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_multi.py -q -m gpu > $out/r02s_pytest_default.log 2>&1; echo "pytest (default) rc=$?" >> $L; tail -2 $out/r02s_pytest_default.log >> $L
for k in 10 100 1000; do for w in 0 1 2; do echo "k=$k HB_GT_WARM=$w" >> $L; HB_GT_WARM=$w python tools/gt_bench.py 10000000 10000 $k 3 mxf4 2>&1 | tail -1 >> $L; done; done
cat $L
