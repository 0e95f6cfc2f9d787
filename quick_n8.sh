for w in c4 c2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --workload $w --steps 5 --warmup 3 > gpurun_out/n8_$w.json 2> gpurun_out/n8_$w.err; echo "rc=$?"; tail -c 600 gpurun_out/n8_$w.err; cat gpurun_out/n8_$w.json | cut -c1-1500
done
