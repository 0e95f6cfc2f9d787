timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "merge" > gpurun_out/pytest_mo.log 2>&1; tail -3 gpurun_out/pytest_mo.log
GLIA_B200_TRACE=1 timeout 900 python profiles/measure_merge_order.py 2>&1 | grep -v "^\[glia_b200\] \(phase\|  \)" | grep "merge_order_bc\|R=" | tee gpurun_out/mo.txt
