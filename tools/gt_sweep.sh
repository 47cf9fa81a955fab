for k in 10 16 32 100 1000; do timeout 150 python tools/gt_bench.py 10000000 10000 $k 2 mxf4,mma 2>&1 | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('k', d['k'], 'mxf4 %.1f ms  mma %.1f ms  identical %s' % (d['mxf4']['ms_best'], d['mma']['ms_best'], d['identical']))"; done
timeout 100 python tools/gt_bench.py 1000000 10000 10 2 mxf4,mma,scan 2>&1 | tail -1 | cut -c1-200
timeout 100 python tools/gt_bench.py 10000000 1000 10 2 mxf4,mma 2>&1 | tail -1 | cut -c1-200
timeout 100 python tools/gt_bench.py 10000000 128 100 2 mxf4,mma 2>&1 | tail -1 | cut -c1-200
