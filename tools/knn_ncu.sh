#!/bin/bash
# one ncu --set full capture of the tensor-core kNN kernel (C2 per-pair shape) -> gpurun_out/knn_mma.ncu-rep
mkdir -p gpurun_out
B="python bench.py --no-cpu-baseline --map-rows 0 --latency-iters 0 --stereo-pairs 0 --search-frames 0 --c4-frames 0 --steps 2 --warmup 1"
$B > gpurun_out/knn_ncu_plain.log 2>&1 || { tail -5 gpurun_out/knn_ncu_plain.log; exit 1; }
ncu --set full --import-source on --clock-control none -k regex:knn2_mma -c 1 -o gpurun_out/knn_mma -f $B > gpurun_out/knn_ncu.log 2>&1
tail -3 gpurun_out/knn_ncu.log
