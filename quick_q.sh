export GLIA_B200_VOXEL=queue
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sharded.py -m gpu -x -q > gpurun_out/pytest_q.log 2>&1; tail -8 gpurun_out/pytest_q.log
for v in queue fast; do export GLIA_B200_VOXEL=$v; for w in c2 c4s; do timeout 200 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bq_${v}_$w.json 2> gpurun_out/bq_${v}_$w.err; tail -c 300 gpurun_out/bq_${v}_$w.err; python -c "
import json,sys; d=json.load(open('gpurun_out/bq_${v}_$w.json')); print('$v $w', round(d['value']), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['stage_ms'].items()}, round(d['roofline']['frac'],4))"; done; done
