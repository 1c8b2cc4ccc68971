#!/bin/bash
# Round 2, GPU call 11 (one B200): sample lanes per pixel for 8 and 4 shares of the thin-lens sphere world (C4) and 2 shares of C2.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for cfg in "8 4 1920 1080 12" "4 4 1920 1080 12" "2 2 1280 720 16" "2 4 1920 1080 12"; do
  for l in -1 3 4 5; do
    echo "== $cfg lanes_log2 $l"; python tools/tile_balance.py $cfg $l | tail -2
  done
done > $O/r02k_lanes_for_shares_spheres.txt 2>&1
cat $O/r02k_lanes_for_shares_spheres.txt
