#!/bin/bash
# Round 2, GPU call 9 (one B200): compute-sanitizer memcheck (one tool) of small renders of every k_render family through the
# host executable; full logs kept.
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
B=pathtracer_b200/bin/pathtracer
: > $O/r02i_memcheck.txt
i=0
for args in "-w1 -p2" "-w2 -p2" "-w3 -p2" "-w3 -p2 -d" "-w4 -p2" "-w5 -p3" "-w5 -p2 -d" "-w5 -p2 --lbvh" "-w2 -p2 --kernel wavefront" "-w5 -p2 --kernel wavefront"; do
  i=$((i+1))
  echo "== pathtracer $args --width 200 --height 120" >> $O/r02i_memcheck.txt
  $B $args --width 200 --height 120 --out $O/mc_plain.bmp > $O/mc_plain_$i.log 2>&1
  echo "plain run exit $?" >> $O/r02i_memcheck.txt
  timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 $B $args --width 200 --height 120 --out $O/mc.bmp > $O/mc_san_$i.log 2>&1
  echo "compute-sanitizer exit $?" >> $O/r02i_memcheck.txt
  grep -E "ERROR SUMMARY|Invalid|LEAK|Error" $O/mc_san_$i.log | head -5 >> $O/r02i_memcheck.txt
done
rm -f $O/mc.bmp $O/mc_plain.bmp
cat $O/r02i_memcheck.txt; tail -5 $O/mc_plain_1.log; tail -8 $O/mc_san_1.log
