set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/s35_gputests.log 2>&1; echo "tests rc=$?" >> gpurun_out/s35_gputests.log
for mode in off last cv; do
  timeout 300 python tools/track_sequence.py --frames 100 --chain $mode >> gpurun_out/s35_chain.jsonl 2>> gpurun_out/s35_chain.err
done
timeout 300 python tools/track_sequence.py --frames 100 --chain cv --trans-cm 3 --rot-deg 1.5 >> gpurun_out/s35_chain.jsonl 2>> gpurun_out/s35_chain.err
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/s35_bench.json 2> gpurun_out/s35_bench.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s35_smoke.log 2>&1
tail -3 gpurun_out/s35_gputests.log; cat gpurun_out/s35_chain.jsonl; tail -3 gpurun_out/s35_chain.err; cat gpurun_out/s35_bench.json; tail -2 gpurun_out/s35_smoke.log
