#!/bin/bash
# A/B of two builds of the library on the GPU box: tools/bin/lib_before.so against the one in place, alternating.
#   bash tools/run_lib_ab.sh [rounds]
L=gsplatloc_b200/lib/libgsplatloc_b200.so
cp $L tools/bin/lib_after.so
for r in $(seq 1 ${1:-2}); do
  for v in after before; do
    cp tools/bin/lib_$v.so $L
    echo "$v replica200k:"; bash tools/quick_bench.sh
    echo "$v tum77k:"; bash tools/quick_bench.sh --workload tum77k
  done
done
cp tools/bin/lib_after.so $L
