set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python profiles/loop_trace.py > gpurun_out/loop_trace.txt 2>&1; tail -12 gpurun_out/loop_trace.txt
timeout 300 python profiles/mapping_bench.py 40 8 > gpurun_out/mapping.txt 2>&1; tail -5 gpurun_out/mapping.txt
B200REG_HOST_LOOP=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_hostloop.csv python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_bench.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:knn1 -c 5 -o gpurun_out/knn1 -f python profiles/prof_knn.py > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
