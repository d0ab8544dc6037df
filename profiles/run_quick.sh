# quick check of a kernel change: k-NN parity tests, the bench line, the device trace of the resident loop
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json
timeout 300 python profiles/loop_trace.py > gpurun_out/loop_trace.txt 2>&1; tail -12 gpurun_out/loop_trace.txt
timeout 300 python profiles/mapping_bench.py 40 8 > gpurun_out/mapping.txt 2>&1; head -2 gpurun_out/mapping.txt; tail -5 gpurun_out/mapping.txt
