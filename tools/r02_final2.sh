#!/bin/bash
# 1-GPU session q: builder change (star rows capped at M): builder + parity tests, bench, inspect of the built top layer.
out=gpurun_out; mkdir -p $out
timeout 1200 python -m pytest tests/test_gpu_builder.py tests/test_gpu_parity.py tests/test_cli.py -q -m gpu > $out/r02q_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $out/r02q_pytest.log
python bench.py > $out/r02q_bench.json 2> $out/r02q_bench.err; echo "bench rc=$?"; grep "^\[bench" $out/r02q_bench.err | tail -3; python -c "
import json
d=json.load(open('$out/r02q_bench.json'))
print('value',d['value'],'frac',d['roofline']['frac'],'kernel_ms',d['roofline']['kernel_ms'],'e2e',d['e2e']['value'],'pinned',d['e2e']['pinned']['value'])
print(d['config']['per_query'], d['config']['recall_at_10'], d['parity'], d['cpu_baseline'])
print(d['strong']['saturating_1M']['qps'], d['strong']['concurrent_callers']['qps'])"
