#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_parity.py -m gpu -q -x > gpurun_out/t5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t5_pytest.log
tail -12 gpurun_out/t5_pytest.log
CMD="python bench.py --workload c4s --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/t5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rf_score_blk -s 2 -c 1 -o gpurun_out/t5_rf $CMD > gpurun_out/t5_ncu.log 2>&1
echo "ncu rc=$?"
