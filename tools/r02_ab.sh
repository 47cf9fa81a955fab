#!/bin/bash
# Round-2 GPU session f: full -m gpu suite, speculation variants (row ahead / sketch prefetch / tail only), latency.
out=gpurun_out; mkdir -p $out
L=$out/r02f_ab.log; : > $L
timeout 1500 python -m pytest tests -q -m gpu > $out/r02f_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?" >> $L; tail -8 $out/r02f_pytest_gpu.log >> $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
for v in cpF s1 s2 s3 s3b s2b s3A; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 2>&1 | tail -1 >> $L; done
for v in s3 s2; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run ${v}_sat 64 10 2>&1 | tail -1 >> $L; done
for ef in 16 128 256; do for v in cpF s3; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v $ef 2>&1 | tail -1 >> $L; done; done
HB_LIB_PATH=$PWD/variants/s3time.so python tools/ab_time.py 64 2>&1 | tail -8 >> $L
cat $L
