#!/bin/bash
# 8-GPU session: the bench under torchrun at N = 8 (weak, split, sweep, sharded ground truth, 100 M sharded HNSW).
out=gpurun_out; mkdir -p $out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > $out/r02p_bench_n8.json 2> $out/r02p_bench_n8.err; echo "bench n8 rc=$?"; grep "^\[bench" $out/r02p_bench_n8.err; grep -v "^frame\|^$\|\*\*\*\|OMP_NUM\|^\[bench" $out/r02p_bench_n8.err | tail -5; head -c 500 $out/r02p_bench_n8.json; echo
