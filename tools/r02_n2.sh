#!/bin/bash
# 2-GPU session: multi-device tests (NCCL inside the library from one process), then the bench under torchrun at N = 2
# with the configs[4] leg forced on smaller shards (the multi-rank code path of the 100 M leg).
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu > $out/r02l_pytest_multi_n2.log 2>&1; echo "pytest multi (2 GPUs) rc=$?"; tail -4 $out/r02l_pytest_multi_n2.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --force-sharded-hnsw --shard-rows 3000000 > $out/r02l_bench_n2.json 2> $out/r02l_bench_n2.err; echo "bench n2 rc=$?"; grep -v "^frame\|^$\|\*\*\*\|OMP_NUM" $out/r02l_bench_n2.err | tail -8; head -c 400 $out/r02l_bench_n2.json; echo; python -c "
import json
d=json.loads([l for l in open('$out/r02l_bench_n2.json') if l.startswith('{')][0])
print(json.dumps(d['strong'].get('sharded_hnsw_100M'), indent=1)); print(d['clocks'])"
