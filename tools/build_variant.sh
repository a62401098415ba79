#!/bin/bash
# A/B build of libsyn3r_warp.so with extra -D flags: tools/build_variant.sh NAME -DSYN3R_FX_ROWS=4 ...
# -> syn3r_b200/_lib/variants/libsyn3r_warp_NAME.so (git-ignored, travels to the GPU box); select it with SYN3R_WARP_LIB.
set -e
name=$1; shift
cd "$(dirname "$0")/../syn3r_b200/csrc"
mkdir -p ../_lib/variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-Wall -Xptxas -v \
    "$@" -shared -o ../_lib/variants/libsyn3r_warp_$name.so abi.cu abi_inverse.cu abi_post.cu 2> ../_lib/variants/$name.ptxas.log
grep -A2 "${GREP_KERNEL:-k_fwd_tile_fxILb1ELb1ELb1}" ../_lib/variants/$name.ptxas.log | grep -E 'registers|spill' | head -4
