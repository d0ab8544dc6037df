# ncu --set full of the five match launches of ONE host-driven registration (lanes per query and seeding as in the graph loop)
mkdir -p gpurun_out
B200REG_HOST_LOOP=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:knn1 -c 5 -o gpurun_out/knn1_hostloop -f python profiles/prof_knn.py > gpurun_out/ncu_full_hostloop.log 2>&1; echo "ncu rc=$?"
