#!/bin/bash
# Round 2, GPU call 7 (EIGHT B200s): the tile-sharded configurations at N = 8 (and the headline at 4) with the diagonal tile interleave.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
export STEPS=20
NS="4 8" tools/scale_run.sh $O/r02g_scale_w5_1080p.jsonl
NS="8" tools/scale_run.sh $O/r02g_scale_c1_w1.jsonl --world 1 --pp 16 --width 1280 --height 720
NS="8" tools/scale_run.sh $O/r02g_scale_c2_w2.jsonl --world 2 --pp 16 --width 1280 --height 720
NS="8" tools/scale_run.sh $O/r02g_scale_c4_w4.jsonl --world 4 --pp 23 --width 1920 --height 1080
NS="8" STEPS=8 tools/scale_run.sh $O/r02g_scale_c5_w5_4k.jsonl --world 5 --pp 32 --width 3840 --height 2160
NS="8" tools/scale_run.sh $O/r02g_scale_c3_w3_tiles.jsonl --world 3 --pp 32 --width 1280 --height 720 --shard tiles
