#!/bin/bash
# First GPU call of the next round: experiments prepared without a GPU at the end of round 1, each checked for
# bit-identity of every output against the product build (tools/ab_identity.py) and timed.  ~25 s of GPU per line.
#   1. -DPT_X_BARE_MUFU      bare MUFU.RCP / MUFU.RSQ instead of __fdividef / rsqrtf (80 instructions off k_render<624>)
#   2. environment sweeps on the product build: PT_SHADE_BATCH x PT_COOP_MAX, PT_TRI_LEAF (the optimum may have moved
#      now that sparse walks are cooperative)
#   3. -DPT_RENDER_MIN_BLOCKS=7 / 6: fewer CTAs per SM for the triangle family (spills 116 -> 52 -> 8 bytes); read w5 only,
#      the flag caps every variant
#   4. -DPT_X_SMEM_STACK: per-lane traversal stack in shared memory (16 KB per CTA) instead of local memory -- local
#      memory is half of the L1 traffic and all of the L2 traffic of k_render (profiles/r01h_k_render_*_summary.txt)
#   6. -DPT_X_PHILOX_KEYS: the ten Philox round keys of the render seed precomputed into the kernel parameters (20 UIADD3 per
#      block become 5 LDCU.128; -40 instructions in k_render<624>)      tools/build_flag_variant.sh philox_keys -DPT_X_PHILOX_KEYS
# Build the variants HERE (CPU container) before the call: tools/r02_build_variants.sh, i.e.
#   tools/build_flag_variant.sh bare_mufu -DPT_X_BARE_MUFU; tools/build_flag_variant.sh mb7 -DPT_RENDER_MIN_BLOCKS=7; tools/build_flag_variant.sh mb6 -DPT_RENDER_MIN_BLOCKS=6
mkdir -p gpurun_out
python tools/ab_identity.py dump pathtracer_b200/libpt_b200.so gpurun_out/ab_prod.npz > gpurun_out/r02_prod.log 2>&1
for v in build/libpt_b200_*.so; do
  [ -e "$v" ] || continue
  t=$(basename $v .so)
  python tools/ab_identity.py dump $v gpurun_out/ab_$t.npz > gpurun_out/r02_$t.log 2>&1
  python tools/ab_identity.py compare gpurun_out/ab_prod.npz gpurun_out/ab_$t.npz | tee gpurun_out/r02_compare_$t.json
  rm -f gpurun_out/ab_$t.npz
done
rm -f gpurun_out/ab_prod.npz
for sb in 12 16 20 24 28 32; do for cm in 8 12 16; do
  echo "shade_batch=$sb coop_max=$cm: $(PT_SHADE_BATCH=$sb PT_COOP_MAX=$cm python tools/profile_render.py --world 5 --pp 16 --frames 3 | tail -1)"
done; done | tee gpurun_out/r02_sweep_shade_coop.txt
for leaf in 2 3 4 6 8; do
  echo "tri_leaf=$leaf: $(PT_TRI_LEAF=$leaf python tools/profile_render.py --world 5 --pp 16 --frames 3 | tail -1)"
done | tee gpurun_out/r02_sweep_tri_leaf.txt
for l in 1 2 3 4 5; do
  echo "sample_lanes_log2=$l: $(PT_SAMPLE_LANES_LOG2=$l python tools/profile_render.py --world 5 --pp 16 --frames 3 | tail -1)"
done | tee gpurun_out/r02_sweep_lanes.txt
# 5. -DPT_X_SHADE_RULE build (tools/build_flag_variant.sh shade_rule -DPT_X_SHADE_RULE): shade the parked lanes also when fewer
#    than PT_SHADE_MIN_TRACE lanes would trace in the next iteration (1 = the product rule "nobody can regenerate")
if [ -e build/libpt_b200_shade_rule.so ]; then
  for mt in 1 4 8 12 16; do for sb in 20 32; do
    echo "min_trace=$mt shade_batch=$sb: $(PT_B200_LIB=build/libpt_b200_shade_rule.so PT_SHADE_MIN_TRACE=$mt PT_SHADE_BATCH=$sb python tools/profile_render.py --world 5 --pp 16 --frames 3 | tail -1)"
  done; done | tee gpurun_out/r02_sweep_shade_rule.txt
fi
