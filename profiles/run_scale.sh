# N-GPU evidence run (one box, N GPUs): the driver's bench line under torchrun (both arms), config #4 sharded over the ranks
# (torchrun + one all_gather) and through the C++ multi-GPU host (one process, ncclCommInitAll).  Usage: bash profiles/run_scale.sh <N> <tag>
N=${1:-2}; TAG=${2:-r2}
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 "${@:2}"; }
timeout 600 bash -c "$(declare -f run); N=$N; run 29511 bench.py --gpus $N --steps 20 --warmup 3" 2> gpurun_out/bench_n${N}_$TAG.err | grep '^{' > gpurun_out/bench_n${N}_$TAG.json; echo "bench n$N rc=$?"
timeout 600 bash -c "$(declare -f run); N=$N; run 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1" 2> gpurun_out/bench_ref_n${N}_$TAG.err | grep '^{' > gpurun_out/bench_ref_n${N}_$TAG.json; echo "ref n$N rc=$?"
timeout 900 bash -c "$(declare -f run); N=$N; run 29513 bench.py --workload config4 --gpus $N --steps 3 --warmup 3 --no-cpu-baseline" 2> gpurun_out/config4_n${N}_$TAG.err | grep '^{' > gpurun_out/config4_n${N}_$TAG.json; echo "config4 n$N rc=$?"
NCCL_DEBUG=INFO timeout 900 python profiles/config4_multi.py 256 200000 3 2> gpurun_out/config4_multi_n${N}_$TAG.err > gpurun_out/config4_multi_n${N}_$TAG.out; echo "multi n$N rc=$?"
grep '^{' gpurun_out/config4_multi_n${N}_$TAG.out > gpurun_out/config4_multi_n${N}_$TAG.json
grep -h -E "NCCL INFO (comm|Connected|Channel 00/|ncclCommInitAll|NVLS)" gpurun_out/config4_multi_n${N}_$TAG.out gpurun_out/config4_multi_n${N}_$TAG.err | head -n 40 > gpurun_out/config4_multi_n${N}_nccl_$TAG.txt
for f in bench_n${N} bench_ref_n${N} config4_n${N} config4_multi_n${N}; do echo "== $f"; cut -c1-600 gpurun_out/${f}_$TAG.json; done
