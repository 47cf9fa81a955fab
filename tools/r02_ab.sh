#!/bin/bash
# Round-2 A/B of the search kernel on the GPU box (gpurun): timings of variants, latency distribution, multi-device tests.
out=gpurun_out; mkdir -p $out
L=$out/r02c_ab.log; : > $L
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_parity.py -x -q -m gpu > $out/r02c_pytest.log 2>&1; echo "pytest rc=$?" >> $L; tail -5 $out/r02c_pytest.log >> $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
python tools/ab.py run v2_default 64 >> $L 2>&1
for v in m7l cpA cpB cpD cpF; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 >> $L 2>&1; done
for v in m7l cpA cpB cpF; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run ${v}_sat 64 10 >> $L 2>&1; done
for ef in 16 32 128 256; do for v in m7l cpA; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v $ef >> $L 2>&1; done; done
HB_LIB_PATH=$PWD/variants/cpTime.so python tools/ab_time.py 64 >> $L 2>&1
cat $L
