#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/t2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t2_pytest.log
tail -15 gpurun_out/t2_pytest.log
CMD="python bench.py --workload c4s --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/t2_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:voxel_pass_fast -s 2 -c 1 -o gpurun_out/t2_voxel $CMD > gpurun_out/t2_ncu.log 2>&1
echo "ncu rc=$?"
