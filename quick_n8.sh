for w in c4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --workload $w --steps 5 --warmup 3 > gpurun_out/n8_$w.json 2> gpurun_out/n8_$w.err; echo "rc=$?"; tail -c 300 gpurun_out/n8_$w.err; python -c "
import json; d=json.load(open('gpurun_out/n8_$w.json')); print(d['value'], d['ms_per_step'], d['stage_ms_rank0'], d['exchange_detail_ms_rank0'], d['e2e'])"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 4 --workload c4 --steps 5 --warmup 3 --no-e2e > gpurun_out/n4_c4.json 2> gpurun_out/n4_c4.err; echo "rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/n4_c4.json')); print(d['value'], d['ms_per_step'], d['stage_ms_rank0'], d['exchange_detail_ms_rank0'])"
