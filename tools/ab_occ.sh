#!/bin/bash
# A/B helper: default build vs a raster_bwd compile-time variant passed in $1 (e.g. -DGSB_BWD_WARPS=2)
bash tools/quick_bench.sh
GSB_NVCC_EXTRA="$1" python gsplatloc_b200/build.py --force > /dev/null
echo "$1:"; bash tools/quick_bench.sh
