export GLIA_B200_VOXEL=queue
timeout 200 python bench.py --workload c4s --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/nq.json 2> gpurun_out/nq.err && \
timeout 600 ncu --set full --import-source on --clock-control none -k regex:voxel_pass_rq -s 3 -c 1 -o gpurun_out/v5b_c4s -f python bench.py --workload c4s --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_q.log 2>&1
tail -3 gpurun_out/ncu_q.log
