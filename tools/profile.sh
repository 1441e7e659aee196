#!/bin/bash
# Profiling pass run on the GPU box (one gpurun call):  tools/profile.sh <tag> [model]
#   1. plain run of the exact command (must exit 0 before anything runs under ncu)
#   2. launch list: every launch of one eager forward with its device time  -> gpurun_out/<tag>_launches.csv
#   3. ncu --set full of the tensor-core / attention / LN launches of one ViT layer and of the
#      memory-bound input/output kernels                                   -> gpurun_out/<tag>_{vit,hbm}.ncu-rep
# Numbers printed under ncu are never bench values; bench.py is run separately.
set -u
TAG=${1:-prof}
MODEL=${2:-standard}
OUT=gpurun_out
mkdir -p $OUT
export MDE_NO_GRAPH=1   # eager launches: one ncu record per kernel, same kernels as the graph replays
CMD="python bench.py --no-extras --steps 1 --warmup 1 --model $MODEL"
$CMD > $OUT/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -20 $OUT/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:tc_gemm2_kernel|vit_attn|layernorm_kernel' -s 91 -c 7 \
    -o $OUT/${TAG}_vit -f $CMD > $OUT/${TAG}_ncu2.log 2>&1
echo "vit capture rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:resize_maps|input_scan_resample_kernel|patchify224_kernel|upsample_concat|conv_halo_kernel|conv_small_kernel|instnorm_kernel|tokens_to_grid|range_finalize|tc_gemm_kernel' -s 0 -c 45 \
    -o $OUT/${TAG}_hbm -f $CMD > $OUT/${TAG}_ncu3.log 2>&1
echo "hbm/conv capture rc=$?"
ls -la $OUT
