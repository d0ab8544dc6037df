# round 2: the evidence run (one GPU).  GPU suite, smoke, both bench arms, ncu launch list of the bench command, ncu --set full of the match
# kernel (-> traffic json), the loop trace, the index build, configs 3 / 4 / 5 with their raw lines, the mapping loop's launch list.
# Usage: bash profiles/run_r2.sh <tag> <commit>
TAG=${1:-r2}; COMMIT=${2:-unknown}
mkdir -p gpurun_out
line() { grep '^{' | tail -n 1; }
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_$TAG.log; tail -n 3 gpurun_out/pytest_$TAG.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone_$TAG.log 2>&1; tail -n 1 gpurun_out/pytest_noalloc_alone_$TAG.log
timeout 600 python __graft_entry__.py smoke > gpurun_out/smoke_$TAG.log 2>&1; tail -n 1 gpurun_out/smoke_$TAG.log
timeout 900 python bench.py 2> gpurun_out/bench_$TAG.err | line > gpurun_out/bench_$TAG.json; echo "bench rc=$?"
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 2> gpurun_out/bench_ref_$TAG.err | line > gpurun_out/bench_ref_$TAG.json; echo "bench ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_$TAG.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --knn-reps 2 > gpurun_out/ncu_bench_$TAG.log 2>&1; echo "ncu list rc=$?"
B200REG_HOST_LOOP=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:knn1 -c 5 -o gpurun_out/knn1_hostloop_$TAG -f python profiles/prof_knn.py > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full rc=$?"
python profiles/ncu_traffic.py gpurun_out/knn1_hostloop_$TAG.ncu-rep gpurun_out/ncu_traffic_$TAG.json $COMMIT
python profiles/ncu_extract.py gpurun_out/knn1_hostloop_$TAG.ncu-rep > gpurun_out/ncu_full_knn1_hostloop_$TAG.txt 2>&1
timeout 300 python profiles/loop_trace.py > gpurun_out/loop_trace_$TAG.txt 2>&1; tail -n 4 gpurun_out/loop_trace_$TAG.txt
timeout 300 python profiles/index_build.py > gpurun_out/index_build_$TAG.txt 2>&1; tail -n 2 gpurun_out/index_build_$TAG.txt | cut -c1-300
timeout 1200 python bench.py --workload config3 2> gpurun_out/config3_n1_$TAG.err | line > gpurun_out/config3_n1_$TAG.json; echo "config3 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_mapping_$TAG.csv python bench.py --workload config3 --scans 24 --no-cpu-baseline > gpurun_out/ncu_mapping_$TAG.log 2>&1; echo "ncu mapping rc=$?"
timeout 1200 python bench.py --workload config4 --steps 3 --warmup 3 2> gpurun_out/config4_n1_$TAG.err | line > gpurun_out/config4_n1_$TAG.json; echo "config4 rc=$?"
timeout 1500 python bench.py --workload config5 2> gpurun_out/config5_n1_$TAG.err | line > gpurun_out/config5_n1_$TAG.json; echo "config5 rc=$?"
for f in bench bench_ref config3_n1 config4_n1 config5_n1; do echo "== $f"; cut -c1-300 gpurun_out/${f}_$TAG.json; done
rm -f gpurun_out/knn1_hostloop_$TAG.ncu-rep.tmp
