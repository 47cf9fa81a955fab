#!/bin/bash
# 8-GPU session: the bench under torchrun at N = 8 (weak, split, sweep, sharded ground truth, 100 M sharded HNSW) and the reference arm.
out=gpurun_out; mkdir -p $out
nvidia-smi -L | wc -l
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 10 --warmup 3 > $out/r02j_bench_n8.json 2> $out/r02j_bench_n8.err; echo "bench n8 rc=$?"; tail -5 $out/r02j_bench_n8.err; head -c 700 $out/r02j_bench_n8.json; echo
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > $out/r02j_ref_n8.json 2> $out/r02j_ref_n8.err; echo "ref n8 rc=$?"; tail -3 $out/r02j_ref_n8.err; head -c 500 $out/r02j_ref_n8.json
