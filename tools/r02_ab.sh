#!/bin/bash
# Round-2 GPU session ae: ground truth, k > 16: per-query histogram of appended rows shared by all work items (HB_GT_HIST=1)
# as a tighter bound on the global k-th distance.  Exactness against the scan first, then timings.
out=gpurun_out; mkdir -p $out
L=$out/r02ae_gt_hist.log; : > $L
run() { HB_GT_HIST=$1 timeout 150 python tools/gt_bench.py $2 $3 $4 ${5:-2} ${6:-mxf4,scan} 2>&1 | tail -1 | python -c "
import sys, json
l = sys.stdin.read().strip()
try:
    d = json.loads(l); print('HB_GT_HIST=$1 n=%d nq=%d k=%d mxf4 ms %s identical_to_scan=%s' % (d['n'], d['nq'], d['k'], ['%.3f' % x for x in d['mxf4']['ms_all']], d.get('identical')))
except Exception:
    print('HB_GT_HIST=$1 n=$2 nq=$3 k=$4 FAILED:', l[-300:])
" >> $L; }
run 1 100001 300 100; run 1 50000 257 1000; run 1 300017 2049 32
if grep -q "FAILED\|=False" $L; then echo "wrong or failed: stopping" >> $L; cat $L; exit 0; fi
HB_GT_HIST=1 timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "ground or truth or gt or plans" > $out/r02ae_pytest_hist.log 2>&1; echo "pytest ground-truth tests (HB_GT_HIST=1) rc=$?" >> $L; tail -1 $out/r02ae_pytest_hist.log >> $L
for h in 0 1; do run $h 10000000 10000 100 3 mxf4; done
for h in 0 1; do run $h 10000000 10000 1000 2 mxf4; done
cat $L
