#!/bin/bash
# 2-GPU check of the committed tree after the hb_knns staging change: the bench under torchrun at N = 2 on a 2 M-row
# index (functional check of the e2e legs on both ranks; the full-size N = 2 / N = 8 lines are profiles/r02h, r02p).
out=gpurun_out; mkdir -p $out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --rows 2000000 --no-extras --no-cpu-baseline > $out/r02ad_bench_n2_2M.json 2> $out/r02ad_bench_n2_2M.err; echo "bench n2 rc=$?"; grep "^\[bench" $out/r02ad_bench_n2_2M.err | tail -3; python -c "
import json
d=json.loads([l for l in open('$out/r02ad_bench_n2_2M.json') if l.startswith('{')][0])
print('value', d['value'], 'e2e', d['e2e']['value'], 'pinned', d['e2e']['pinned']['value'], 'parity', d.get('parity'), 'gt', d['ground_truth_sharded']['ms'], d['ground_truth_sharded'].get('checksum'))"
