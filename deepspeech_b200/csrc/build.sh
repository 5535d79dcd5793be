#!/bin/bash
# Builds libds2_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -Wall --use_fast_math -Xptxas -v"
mkdir -p build
PIDS=()
for f in ds2_api ds2_ctc ds2_generic ds2_rnn_generic ds2_gemm ds2_rnn_persistent ds2_rnn_persist2 ds2_gru_persistent ds2_conv_gemm ds2_conv_window ds2_conv1_fused ds2_optim ds2_decode; do
  stale=0; for dep in $f.cu *.cuh ../../include/ds2_b200.h build.sh; do [ "$dep" -nt build/$f.o ] && stale=1; done
  if [ ! -f build/$f.o ] || [ $stale = 1 ]; then
    ( $NVCC $FLAGS -c $f.cu -o build/$f.o > build/$f.log 2>&1 || { cat build/$f.log; exit 1; } ) &
    PIDS+=($!)
  fi
done
for p in "${PIDS[@]}"; do wait $p; done
$NVCC -gencode arch=compute_100a,code=sm_100a -shared -o libds2_b200.so build/*.o -cudart static
echo "built $(pwd)/libds2_b200.so"
