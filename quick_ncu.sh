timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/nq.json 2> gpurun_out/nq.err && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1_launches_c2.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_l.log 2>&1
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'voxel_pass_fast|rf_score_smem|init_candidates_blocks|gather_edges' -s 12 -c 4 -o gpurun_out/r1_end_c2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_q.log 2>&1
tail -3 gpurun_out/ncu_q.log
