#!/bin/bash
# Experimental build of the library with extra nvcc flags into build/exp/<tag>/libavdiff_b200.so (load it with AVD_LIB_PATH).
#   usage: tools/exp_build.sh <tag> [nvcc flags ...]      e.g. tools/exp_build.sh nowait -DAVD_EXP_NOWAIT
set -e
tag=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
out=$root/build/exp/$tag; mkdir -p $out
cs=$root/multimodal_diffusion_b200/csrc
pids=()
for f in avd_elementwise avd_gemm avd_gemm_ps avd_gemm_chain avd_gemm_ps2 avd_attn avd_attn_fa avd_stream avd_codec avd_vae avd_api; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC "$@" -c $cs/$f.cu -o $out/$f.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o $out/libavdiff_b200.so $out/*.o -gencode arch=compute_100a,code=sm_100a
echo $out/libavdiff_b200.so
