#!/bin/bash
# Round-2 GPU session y: ground-truth kernel plans (HB_GT_PLAN 0 = 240-row tiles x 2 accumulators, 1 = CTA pair / cta_group::2,
# 2 = 120-row tiles x 4 accumulators): exactness against the scan on ragged sizes, then the 10 M x 10 k timings.
out=gpurun_out; mkdir -p $out
L=$out/r02y_gt_plans.log; : > $L
run() { HB_GT_PLAN=$1 timeout 150 python tools/gt_bench.py $2 $3 $4 ${5:-2} ${6:-mxf4,scan} 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    d = json.loads(l); print('HB_GT_PLAN=$1 n=%d nq=%d k=%d mxf4 ms %s identical_to_scan=%s' % (d['n'], d['nq'], d['k'], ['%.3f' % x for x in d['mxf4']['ms_all']], d.get('identical')))
except Exception:
    print('HB_GT_PLAN=$1 n=$2 nq=$3 k=$4 FAILED:', l[-300:])
" >> $L; }
for pl in 0 1 2; do run $pl 300017 2049 10; run $pl 100001 300 100; run $pl 50000 257 1000; done
for k in 10 100; do for pl in 0 1 2; do run $pl 10000000 10000 $k 3 mxf4; done; done
HB_GT_PLAN=1 timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "ground or truth or gt or tied or smoke or cross" > $out/r02y_pytest_plan1.log 2>&1; echo "pytest ground-truth tests (HB_GT_PLAN=1) rc=$?" >> $L; tail -1 $out/r02y_pytest_plan1.log >> $L
cat $L
