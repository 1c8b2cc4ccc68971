#!/bin/bash
# Round 2, GPU call 8 (one B200): the GPU suite on the final tree, then compute-sanitizer memcheck (one tool) of small renders of
# every world through the host executable (every k_render family, the exchange kernels via --gpus 1 are not reached: see the tests).
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
( timeout 1200 python -m pytest tests -m gpu -q > $O/r02h_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $O/r02h_pytest_gpu.txt )
tail -4 $O/r02h_pytest_gpu.txt
B=pathtracer_b200/bin/pathtracer
: > $O/r02h_memcheck.txt
for args in "-w1 -p2" "-w2 -p2" "-w3 -p2" "-w3 -p2 -d" "-w4 -p2" "-w5 -p3" "-w5 -p2 -d" "-w5 -p2 --lbvh" "-w2 -p2 --kernel wavefront" "-w5 -p2 --kernel wavefront"; do
  echo "== pathtracer $args --width 200 --height 120" >> $O/r02h_memcheck.txt
  $B $args --width 200 --height 120 --out $O/mc_plain.bmp > /dev/null 2>&1 &&
  timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 $B $args --width 200 --height 120 --out $O/mc.bmp 2>&1 | grep -E "ERROR SUMMARY|Invalid|Error|error" | head -5 >> $O/r02h_memcheck.txt
done
rm -f $O/mc.bmp $O/mc_plain.bmp
cat $O/r02h_memcheck.txt
