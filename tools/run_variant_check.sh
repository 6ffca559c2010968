#!/bin/bash
# SURVEY.md Appendix B switches, kernel side against oracle side (GPU box): build the library with the alternative constants
# and hold it to the oracle set the same way; then restore the default build.
GSB_NVCC_EXTRA="-DGSB_RADIUS_FLOOR=0.1f -DGSB_CLAMP_ASYMMETRIC=1" python gsplatloc_b200/build.py --force > /dev/null
GSB_ORACLE_RADIUS_FLOOR=0.1 GSB_ORACLE_CLAMP=asymmetric timeout 600 python -m pytest -q -m gpu \
  tests/test_gpu_tracker.py::test_one_iteration_matches_oracle tests/test_gpu_tracker.py::test_intermediates_bit_exact_given_the_kernel_projection \
  "tests/test_gpu_parity_full.py::test_one_iteration_at_benchmark_size_matches_oracle[tum77k]" 2>&1 | tail -2
python gsplatloc_b200/build.py --force > /dev/null
