#!/bin/bash
# Round-2 evidence run on one B200: GPU tests, bench lines (headline, TUM shape, reference arm), then -- after those exited
# 0 without ncu -- the ncu launch list and one full capture of the iteration's kernels.  Everything lands in gpurun_out/.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_gputests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2_gputests.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err
timeout 300 python bench.py --steps 20 --warmup 3 --workload tum77k --no-cpu-baseline > gpurun_out/r2_bench_tum77k.json 2>> gpurun_out/r2_bench.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2>> gpurun_out/r2_bench.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2_smoke.log 2>&1
python tools/prof_iter.py --advance 20 --iters 3 > gpurun_out/r2_plain.log 2>&1 &&
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv python tools/prof_iter.py --advance 20 --iters 3 > gpurun_out/r2_ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"raster|tile_sort|project_emit|loss|pose_finish" -c 12 -o gpurun_out/r2_iter python tools/prof_iter.py --advance 20 --iters 2 > gpurun_out/r2_ncu_f.log 2>&1
tail -3 gpurun_out/r2_gputests.log; cat gpurun_out/r2_bench.json | cut -c1-400; tail -2 gpurun_out/r2_smoke.log
