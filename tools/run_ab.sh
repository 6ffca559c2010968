#!/bin/bash
# Round-2 A/B helper (GPU box): GPU tests, then quick benches of kernel variants, then one ncu capture of the rasterisers.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/ab_tests.log
for v in "1 1" "1 2" "2 2"; do
  set -- $v
  echo "fwd=$1 bwd=$2:"; GSB_FWD_VARIANT=$1 GSB_BWD_VARIANT=$2 bash tools/quick_bench.sh ${AB_ARGS}
done
if [ -n "$AB_NCU" ]; then
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$AB_NCU" -c ${AB_NCU_COUNT:-4} -o gpurun_out/${AB_NAME:-ab}_prof python tools/prof_iter.py --advance 20 --iters 2 > gpurun_out/ab_ncu.log 2>&1; tail -2 gpurun_out/ab_ncu.log
fi
