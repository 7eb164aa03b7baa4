#!/bin/bash
# developer helper: tools/build_variant.sh <name> [extra nvcc flags...]  ->  build/lib_<name>.so  (for tools/ab.sh)
set -e
name="$1"; shift
here="$(cd "$(dirname "$0")/.." && pwd)"
mkdir -p "$here/build"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -shared "$@" \
     -o "$here/build/lib_${name}.so" "$here/bayesian_inference_b200/csrc/bi_api.cu"
echo "built build/lib_${name}.so"
