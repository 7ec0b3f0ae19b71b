#!/bin/bash
# queue-dry / last-exit times of a C2-geometry launch by item count (RTB200_SPLITS_LOG2), -DRTB_TIMELINE build
mkdir -p gpurun_out
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
cp build_ab/librtb200_timeline.so rust-ray-tracing_b200/librtb200.so
for spp in 100 125; do for lg in 1 2 3 4 5; do
  echo "spp $spp lg $lg: $(RTB200_SPLITS_LOG2=$lg python tools/timeline_run.py $spp 2>&1 | sed -n 1,2p | tr '\n' ' ')"
done; done | tee gpurun_out/timeline_k.txt
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
