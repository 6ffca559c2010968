#!/bin/bash
# A/B helper (GPU box): bench the library in place, then rebuild with each compile-time variant in "$@" and bench again.
export GSB_FWD_VARIANT=${GSB_FWD_VARIANT:-1}
echo "default:"; bash tools/quick_bench.sh
for v in "$@"; do
  GSB_NVCC_EXTRA="$v" python gsplatloc_b200/build.py --force > /dev/null
  echo "$v:"; bash tools/quick_bench.sh
done
