#!/bin/bash
# BASELINE.json configs C1..C5 (+ the 1080p headline) at N GPUs (NS, default "1 2 4 8") -> <prefix>_<cfg>.jsonl
p=${1:-gpurun_out/scale}
export STEPS=${STEPS:-20}
tools/scale_run.sh ${p}_c1_w1.jsonl --world 1 --pp 16 --width 1280 --height 720
tools/scale_run.sh ${p}_c2_w2.jsonl --world 2 --pp 16 --width 1280 --height 720
tools/scale_run.sh ${p}_c3_w3_samples.jsonl --world 3 --pp 32 --width 1280 --height 720 --shard samples
tools/scale_run.sh ${p}_c4_w4.jsonl --world 4 --pp 23 --width 1920 --height 1080
STEPS=8 tools/scale_run.sh ${p}_c5_w5_4k.jsonl --world 5 --pp 32 --width 3840 --height 2160
tools/scale_run.sh ${p}_w5_1080p.jsonl
