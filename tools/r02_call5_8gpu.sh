#!/bin/bash
# Round 2, GPU call 5 (EIGHT B200s): BASELINE configs C1..C5 and the 1080p headline at N = 2, 4, 8 ranks (N = 1 lines come
# from the one-GPU call, profiles/r02c_bench_configs_1gpu.jsonl; the headline's N = 1 is repeated here), >= 20 timed steps,
# clocks sampled, per-rank phase times in every line.  tiles_p2p next to the NCCL gather on the two smallest frames.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/r02e_topo.txt 2>&1
export STEPS=20
NS="1 2 4 8" tools/scale_run.sh $O/r02e_scale_w5_1080p.jsonl
NS="4 8" tools/scale_run.sh $O/r02e_scale_c1_w1.jsonl --world 1 --pp 16 --width 1280 --height 720
NS="2 4 8" tools/scale_run.sh $O/r02e_scale_c2_w2.jsonl --world 2 --pp 16 --width 1280 --height 720
NS="4 8" tools/scale_run.sh $O/r02e_scale_c3_w3_samples.jsonl --world 3 --pp 32 --width 1280 --height 720 --shard samples
NS="2 4 8" tools/scale_run.sh $O/r02e_scale_c4_w4.jsonl --world 4 --pp 23 --width 1920 --height 1080
NS="2 4 8" STEPS=8 tools/scale_run.sh $O/r02e_scale_c5_w5_4k.jsonl --world 5 --pp 32 --width 3840 --height 2160
NS="8" tools/scale_run.sh $O/r02e_scale_c1_w1_tiles_p2p.jsonl --world 1 --pp 16 --width 1280 --height 720 --shard tiles_p2p
NS="8" tools/scale_run.sh $O/r02e_scale_c2_w2_tiles_p2p.jsonl --world 2 --pp 16 --width 1280 --height 720 --shard tiles_p2p
NS="8" tools/scale_run.sh $O/r02e_scale_c3_w3_tiles.jsonl --world 3 --pp 32 --width 1280 --height 720 --shard tiles
