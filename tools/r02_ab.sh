#!/bin/bash
# Round-2 GPU session: CLI tests, A/B of gather variants, latency distribution, reference-built index comparison, bench.
out=gpurun_out; mkdir -p $out
L=$out/r02d_ab.log; : > $L
timeout 600 python -m pytest tests/test_cli.py -x -q -m gpu > $out/r02d_pytest_cli.log 2>&1; echo "pytest cli rc=$?" >> $L; tail -5 $out/r02d_pytest_cli.log >> $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
python tools/ab.py run v2_default 64 >> $L 2>&1
for v in m7l m8l cpA cpB cpD cpF; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v 64 2>&1 | tail -1 >> $L; done
for v in m7l cpA cpB cpF; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run ${v}_sat 64 10 2>&1 | tail -1 >> $L; done
for ef in 16 32 128 256; do for v in m7l cpA; do HB_LIB_PATH=$PWD/variants/$v.so python tools/ab.py run $v $ef 2>&1 | tail -1 >> $L; done; done
HB_LIB_PATH=$PWD/variants/cpTime.so python tools/ab_time.py 64 2>&1 | tail -8 >> $L
cat $L
python tools/refbuilt.py 1000000 10000 > $out/r02d_refbuilt_1M.json 2> $out/r02d_refbuilt.err; echo "refbuilt rc=$?"; tail -3 $out/r02d_refbuilt.err
python bench.py --steps 10 --warmup 3 > $out/r02d_bench.json 2> $out/r02d_bench.err; echo "bench rc=$?"; tail -5 $out/r02d_bench.err; cat $out/r02d_bench.json | head -c 3000
