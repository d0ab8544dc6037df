# round 2: the evidence run.  GPU suite, bench line, ncu launch list of the bench command, ncu --set full of the match kernel (-> traffic json),
# the loop trace, the mapping loop with its launch list.  Usage: bash profiles/run_r2.sh <tag> <commit>
TAG=${1:-r2}; COMMIT=${2:-unknown}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_$TAG.log; tail -n 3 gpurun_out/pytest_$TAG.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -n 1 gpurun_out/smoke_$TAG.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --knn-reps 2 > gpurun_out/ncu_bench_$TAG.log 2>&1; echo "ncu list rc=$?"
B200REG_HOST_LOOP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:knn1 -c 5 -o gpurun_out/knn1_hostloop_$TAG -f python profiles/prof_knn.py > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
python profiles/ncu_traffic.py gpurun_out/knn1_hostloop_$TAG.ncu-rep gpurun_out/ncu_traffic_$TAG.json $COMMIT
python profiles/ncu_extract.py gpurun_out/knn1_hostloop_$TAG.ncu-rep > gpurun_out/ncu_full_knn1_hostloop_$TAG.txt 2>&1
timeout 300 python profiles/loop_trace.py > gpurun_out/loop_trace_$TAG.txt 2>&1; tail -n 8 gpurun_out/loop_trace_$TAG.txt
timeout 900 python bench.py --workload config3 --scans 100 > gpurun_out/config3_100_$TAG.json 2> gpurun_out/config3_100_$TAG.err; echo "config3 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mapping_$TAG.csv python bench.py --workload config3 --scans 24 --no-cpu-baseline > gpurun_out/ncu_mapping_$TAG.log 2>&1; echo "ncu mapping rc=$?"
