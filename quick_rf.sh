run() { for w in c2 c4s c5; do timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/brf_$w.json 2> gpurun_out/brf_$w.err; tail -c 300 gpurun_out/brf_$w.err; python -c "
import json,sys; d=json.load(open('gpurun_out/brf_$w.json')); print('$1 $w', round(d['value']), round(d['ms_per_step'],3), d.get('stage_ms',{}).get('rf_score'))"; done; }
run auto
