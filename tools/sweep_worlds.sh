#!/bin/bash
# usage: tools/sweep_worlds.sh "<profile_render.py args>" ...  -- Gsamples/s of the five worlds (BASELINE configs, reduced spp) per spec
for spec in "$@"; do
  echo "== $spec"
  for cfg in "1 1280 720 16" "2 1280 720 16" "3 1280 720 16" "4 1920 1080 12" "5 1920 1080 16"; do
    set -- $cfg
    python tools/profile_render.py --world $1 --width $2 --height $3 --pp $4 --frames 3 --json $spec | tail -1
  done
done
