#!/bin/bash
# 1-GPU session m: full -m gpu suite, the same parity tests on the bounds-checked build, the default bench.
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -q -m gpu > $out/r02m_pytest_gpu.log 2>&1; echo "pytest gpu rc=$?"; tail -4 $out/r02m_pytest_gpu.log
HB_LIB_PATH=$PWD/variants/dbg.so timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_builder.py -q -m gpu > $out/r02m_pytest_debug_bounds.log 2>&1; echo "pytest (bounds-checked build) rc=$?"; tail -4 $out/r02m_pytest_debug_bounds.log
python bench.py > $out/r02m_bench.json 2> $out/r02m_bench.err; echo "bench rc=$?"; grep "^\[bench" $out/r02m_bench.err; head -c 400 $out/r02m_bench.json; echo
python bench.py --impl reference --steps 3 --warmup 1 > $out/r02m_ref.json 2> $out/r02m_ref.err; echo "ref rc=$?"; head -c 300 $out/r02m_ref.json
