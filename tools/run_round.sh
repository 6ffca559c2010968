set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s36_gputests.log 2>&1; echo "tests rc=$?" >> gpurun_out/s36_gputests.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/s36_bench.json 2> gpurun_out/s36_bench.err
timeout 300 python bench.py --steps 20 --warmup 3 --workload tum77k --no-cpu-baseline > gpurun_out/s36_bench_tum.json 2> gpurun_out/s36_bench_tum.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/s36_bench_ref.json 2> gpurun_out/s36_bench_ref.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/s36_launches.csv python tools/prof_iter.py --advance 20 --iters 3 > gpurun_out/s36_ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"raster|tile_sort|project_emit|loss|pose_finish" -c 14 -o gpurun_out/s36_iter python tools/prof_iter.py --advance 20 --iters 2 > gpurun_out/s36_ncu_f.log 2>&1
ls -la gpurun_out
