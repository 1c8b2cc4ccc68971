#!/bin/bash
# Round 2, GPU call 12 (EIGHT B200s): the sphere-tree configurations (C2, C4) at N = 4 and 8 with 8 sample lanes per pixel at every share size.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
export STEPS=20
NS="4 8" tools/scale_run.sh $O/r02l_scale_c4_w4.jsonl --world 4 --pp 23 --width 1920 --height 1080
NS="4 8" tools/scale_run.sh $O/r02l_scale_c2_w2.jsonl --world 2 --pp 16 --width 1280 --height 720
