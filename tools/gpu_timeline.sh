#!/bin/bash
# swaps the -DRTB_TIMELINE build in (build_ab/librtb200_timeline.so) and prints the warp start / dry / exit timeline of a C2 launch
mkdir -p gpurun_out
cp rust-ray-tracing_b200/librtb200.so /tmp/librtb200_keep.so
cp build_ab/librtb200_timeline.so rust-ray-tracing_b200/librtb200.so
python tools/timeline_run.py 100 > gpurun_out/timeline_100.txt 2>&1
python tools/timeline_run.py 25 > gpurun_out/timeline_25.txt 2>&1
cp /tmp/librtb200_keep.so rust-ray-tracing_b200/librtb200.so
head -50 gpurun_out/timeline_100.txt
