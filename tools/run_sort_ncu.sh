#!/bin/bash
# warp instructions / duration / shared wavefronts of tile_sort_fixed_kernel for the library in place and lib_before.so
L=gsplatloc_b200/lib/libgsplatloc_b200.so
cp $L tools/bin/lib_after.so
python tools/prof_iter.py --advance 20 --iters 2 > /dev/null 2>&1 || exit 1
for v in after before; do
  cp tools/bin/lib_$v.so $L
  timeout 300 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active \
    --clock-control none -k regex:tile_sort -c 2 --csv --log-file gpurun_out/sort_ncu_$v.csv python tools/prof_iter.py --advance 20 --iters 2 > gpurun_out/sort_ncu_$v.log 2>&1
  echo "$v:"; grep -v "^==" gpurun_out/sort_ncu_$v.csv | awk -F'","' '{print $5, $(NF-2), $(NF-1), $NF}' | tail -14
done
cp tools/bin/lib_after.so $L
