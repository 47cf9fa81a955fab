#!/bin/bash
# 2-GPU session: multi-device tests (NCCL inside the library from one process), then the bench under torchrun at N = 2.
out=gpurun_out; mkdir -p $out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu > $out/r02h_pytest_multi_n2.log 2>&1; echo "pytest multi (2 GPUs) rc=$?"; tail -6 $out/r02h_pytest_multi_n2.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > $out/r02h_bench_n2.json 2> $out/r02h_bench_n2.err; echo "bench n2 rc=$?"; tail -5 $out/r02h_bench_n2.err; head -c 1500 $out/r02h_bench_n2.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $out/r02h_ref_n2.json 2> $out/r02h_ref_n2.err; echo "ref n2 rc=$?"; tail -3 $out/r02h_ref_n2.err; head -c 600 $out/r02h_ref_n2.json
