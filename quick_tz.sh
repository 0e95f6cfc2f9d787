for tz in 16 32 64 128; do GLIA_B200_TZ=$tz timeout 200 python bench.py --workload c4s --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err; python -c "
import json; d=json.load(open('gpurun_out/b.json')); print('c4s TZ=$tz', round(d['stage_ms']['voxel_pass'],3))"; GLIA_B200_TZ=$tz timeout 200 python bench.py --workload c2 --steps 5 --warmup 2 --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err; python -c "
import json; d=json.load(open('gpurun_out/b.json')); print('c2 TZ=$tz', round(d['stage_ms']['voxel_pass'],3))"; done
