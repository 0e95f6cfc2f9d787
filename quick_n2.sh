timeout 600 python -m pytest tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/pytest_sh.log 2>&1; tail -2 gpurun_out/pytest_sh.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload c4 --steps 5 --warmup 3 --no-e2e > gpurun_out/n2_c4.json 2> gpurun_out/n2_c4.err; echo "rc=$?"; tail -c 400 gpurun_out/n2_c4.err; python -c "
import json; d=json.load(open('gpurun_out/n2_c4.json')); print(d['value'], d['ms_per_step'], d['stage_ms_rank0'], d['exchange_detail_ms_rank0'])"
