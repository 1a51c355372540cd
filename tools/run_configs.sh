#!/bin/bash
# BASELINE.json configs 3 and 4 plus the POPC denominator; one JSON line each into $1 (default gpurun_out/configs.jsonl)
OUT=${1:-gpurun_out/configs.jsonl}; : > $OUT
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/popc_bench tools/popc_bench.cu && /tmp/popc_bench >> $OUT
K="python tools/kernel_times.py"
# config 3: 1920x1080, 8000 keypoints/frame, matcher-bound
for bits in 512 32; do
  $K --tag cfg3_1080p_N8000 --rows 1080 --cols 1920 --max-kp 8000 --threshold 6000 --bits $bits --frames 64 --batch 16 --passes 3 >> $OUT
done
$K --tag cfg3_1080p_N8000_thr20000 --rows 1080 --cols 1920 --max-kp 8000 --bits 512 --frames 64 --batch 16 --passes 3 >> $OUT
# config 4: 3840x2160 stencil sweep over batch size
for b in 1 2 4 8 16 32 64; do
  $K --tag cfg4_4k_batch$b --rows 2160 --cols 3840 --frames 64 --batch $b --passes 2 >> $OUT
done
# config 2 for the record, REF32 and FULL512
$K --tag cfg2_kitti --frames 1000 --batch 250 --bits 32 >> $OUT
$K --tag cfg2_kitti_full512 --frames 1000 --batch 250 --bits 512 >> $OUT
cat $OUT | cut -c1-700
