#!/bin/bash
# A/B helper: default build vs raster_bwd compiled for 3 blocks/SM
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
bash tools/quick_bench.sh
GSB_NVCC_EXTRA=-DGSB_BWD_OCC=3 python gsplatloc_b200/build.py --force > /dev/null
echo "OCC=3:"; bash tools/quick_bench.sh
