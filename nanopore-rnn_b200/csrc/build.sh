#!/usr/bin/env bash
# Builds libnanotensor_b200.so for sm_100a (in-tree, next to the Python package).
set -euo pipefail
# NTB_ABLATION=1 compiles the A/B measurement switches in (ntb_env(), ntb_common.cuh); the default build ignores them.
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
out="$here/../libnanotensor_b200.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
flags=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17
       -Xcompiler -fPIC,-O2,-Wall,-fvisibility=hidden -shared -cudart static)
"$NVCC" "${flags[@]}" ${NTB_PTXAS_V:+-Xptxas -v} ${NTB_DEBUG_SYNC:+-DNTB_DEBUG_SYNC} ${NTB_ABLATION:+-DNTB_ABLATION} -o "$out" "$here/ntb_api.cu" "$here/ntb_gemm_tc.cu"
echo "built $out"
