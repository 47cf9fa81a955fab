#!/bin/bash
# Runs on the GPU box (under gpurun): the plain bench first, then the ncu launch list of the same command
# (index-construction kernels filtered out by name: ~350 k launches that are outside the timed region), one
# `--set full` capture of the search kernel and one of the tcgen05 ground-truth kernel (tools/gt_bench.py).
# Outputs land in gpurun_out/; summaries are copied to profiles/ by hand (tools/ncu_summary.py).
set -u
tag="${1:-r02}"
shift || true
cmd="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras $*"
out=gpurun_out
mkdir -p $out
$cmd > $out/bench_plain_$tag.json 2> $out/bench_plain_$tag.err || { echo "plain run failed"; tail -5 $out/bench_plain_$tag.err; exit 1; }
head -c 600 $out/bench_plain_$tag.json; echo
ncu --metrics gpu__time_duration.sum --clock-control none -k 'regex:hnsw_search|scan_topk|topk_merge|unpack|synth|gt_|tensor_peak|pack_kernel' -c 400 \
    --csv --log-file $out/launches_$tag.csv $cmd > $out/ncu_launches_$tag.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:hnsw_search_v2' -s 4 -c 2 -f -o $out/search_$tag $cmd > $out/ncu_full_$tag.log 2>&1
echo "full capture (search) rc=$?"
gcmd="python tools/gt_bench.py 10000000 10000 10 1 mxf4"
$gcmd > $out/gt_plain_$tag.json 2>&1 && ncu --set full --clock-control none --import-source on -k 'regex:gt_mxf4' -s 1 -c 1 -f -o $out/gt_$tag $gcmd > $out/ncu_gt_$tag.log 2>&1
echo "full capture (ground truth) rc=$?"
ls -la $out | tail -12
