#!/bin/bash
# Builds monogs_b200/lib/libmonogs_b200.so for sm_100a (what __graft_entry__.build() runs).
set -e
cd "$(dirname "$0")"
mkdir -p ../lib
SRCS="raster_forward.cu blend_forward.cu blend_track.cu raster_backward.cu pose_backward.cu loss_adam.cu knn_pose.cu track_fused.cu map_fused.cu densify.cu frame_ops.cu peer_exchange.cu api.cu"
OBJS=""
pids=()
for f in $SRCS; do
  o="../lib/${f%.cu}.o"
  OBJS="$OBJS $o"
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 --expt-relaxed-constexpr \
       -Xcompiler -fPIC ${MGS_NVCC_EXTRA:-} -c -o "$o" "$f" &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../lib/libmonogs_b200.so $OBJS
rm -f $OBJS
# stand-alone checker of the block bitonic sort (tests/test_ops_gpu.py runs it on the GPU box)
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o ../../scripts/tools/bitonic_check ../../scripts/tools/bitonic_check.cu
