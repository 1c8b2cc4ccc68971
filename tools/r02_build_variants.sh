#!/bin/bash
# Builds every experiment variant tools/r02_experiments.sh A/Bs (CPU container, ~40 s each) into build/.
set -e
cd "$(dirname "$0")/.."
tools/build_flag_variant.sh bare_mufu -DPT_X_BARE_MUFU
tools/build_flag_variant.sh philox_keys -DPT_X_PHILOX_KEYS
tools/build_flag_variant.sh diet -DPT_X_BARE_MUFU -DPT_X_PHILOX_KEYS
tools/build_flag_variant.sh mb7 -DPT_RENDER_MIN_BLOCKS=7
tools/build_flag_variant.sh mb6 -DPT_RENDER_MIN_BLOCKS=6
tools/build_flag_variant.sh smem_stack -DPT_X_SMEM_STACK
tools/build_flag_variant.sh shade_rule -DPT_X_SHADE_RULE
