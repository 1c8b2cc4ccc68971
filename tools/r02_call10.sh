#!/bin/bash
# Round 2, GPU call 10 (one B200): sample lanes per pixel for the small 8-GPU shares (launch tail), and the drop-in test with the rebuilt reference binary.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
( timeout 600 python -m pytest tests/test_tools.py -m gpu -q > $O/r02j_pytest_tools.txt 2>&1; echo "pytest exit $?" >> $O/r02j_pytest_tools.txt ); tail -3 $O/r02j_pytest_tools.txt
for cfg in "8 1 1280 720 16" "8 2 1280 720 16" "8 3 1280 720 32" "8 5 1920 1080 32" "4 1 1280 720 16" "4 2 1280 720 16"; do
  for l in -1 3 4 5; do
    echo "== $cfg lanes_log2 $l"; python tools/tile_balance.py $cfg $l | tail -2
  done
done > $O/r02j_lanes_for_shares.txt 2>&1
cat $O/r02j_lanes_for_shares.txt
