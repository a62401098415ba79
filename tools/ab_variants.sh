#!/bin/bash
# Times every A/B build under syn3r_b200/_lib/variants (tools/build_variant.sh) on one workload, with output checksums.
# Run under gpurun: tools/ab_variants.sh [config] [reps] -> gpurun_out/ab_<config>.jsonl
cfg=${1:-B_svd}; reps=${2:-10}
mkdir -p gpurun_out
out=gpurun_out/ab_${cfg}.jsonl; : > $out
for so in syn3r_b200/_lib/libsyn3r_warp.so syn3r_b200/_lib/variants/*.so; do
    SYN3R_WARP_LIB=$PWD/$so SYN3R_TIME_CHECK=1 timeout 300 python tools/time_kernels.py $cfg $reps >> $out 2>> gpurun_out/ab_${cfg}.err || echo "{\"lib\": \"$so\", \"failed\": true}" >> $out
done
cat $out
