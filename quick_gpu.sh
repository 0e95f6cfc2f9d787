timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; tail -8 gpurun_out/pytest_gpu.log
for w in c2 c4; do timeout 200 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 400 gpurun_out/bench_$w.err; python -c "
import json,sys; d=json.load(open('gpurun_out/bench_$w.json')); print('$w', round(d['value']), round(d['ms_per_step'],3), {k:round(v,3) for k,v in d['stage_ms'].items()}, round(d['roofline']['frac'],4))"; done
