#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_layer_kernel_gpu.py -m gpu -q -p no:cacheprovider --timeout=120 --timeout-method=thread 2>&1 | tail -3
timeout 120 python scripts/layer_timeline.py > gpurun_out/r02_layer_timeline_fused.txt 2>&1; cat gpurun_out/r02_layer_timeline_fused.txt | head -60
LIFTER_FUSE_LAYER=0 timeout 120 python scripts/layer_timeline.py > gpurun_out/r02_layer_timeline_plain.txt 2>&1; head -30 gpurun_out/r02_layer_timeline_plain.txt
