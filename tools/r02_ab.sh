#!/bin/bash
# Round-2 GPU session ac: threads per host copy in hb_knns's staged path (HB_HOST_COPY_THREADS, default 4).
out=gpurun_out; mkdir -p $out
L=$out/r02ac_host_copy_threads.log; : > $L
python tools/ab.py prepare 10000000 10000 >> $L 2>&1
for rep in 1 2; do for t in 1 2 4 6 8 12; do echo -n "threads=$t " >> $L; HB_HOST_COPY_THREADS=$t python tools/ab_e2e.py t$t 1 2>&1 | tail -1 >> $L; done; done
for t in 4 8; do echo -n "threads=$t " >> $L; HB_HOST_COPY_THREADS=$t python tools/ab_e2e.py t$t 10 2>&1 | tail -1 >> $L; done
cat $L
