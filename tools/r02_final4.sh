#!/bin/bash
# Final 1-GPU validation of the tree as committed: full -m gpu suite, smoke(), both bench arms.
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -x -q -m gpu > $out/r02ab_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -3 $out/r02ab_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/r02ab_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $out/r02ab_smoke.log
python bench.py --impl reference > $out/r02ab_ref.json 2> $out/r02ab_ref.err; echo "ref rc=$?"; head -c 250 $out/r02ab_ref.json; echo
python bench.py > $out/r02ab_bench.json 2> $out/r02ab_bench.err; echo "bench rc=$?"; grep "^\[bench" $out/r02ab_bench.err | tail -2; python -c "
import json
d=json.load(open('$out/r02ab_bench.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'],'cpu',d['cpu_baseline']['value'],d['cpu_baseline']['cores'],'parity',d['parity'],'gt',d['ground_truth_sharded']['ms'],d['ground_truth_sharded']['roofline']['frac'],d['clocks'])"
