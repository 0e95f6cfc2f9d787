run() { for w in c2 c4s; do timeout 200 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/b.json 2> gpurun_out/b.err; tail -c 200 gpurun_out/b.err; python -c "
import json; d=json.load(open('gpurun_out/b.json')); print('$1 $w', round(d['stage_ms']['voxel_pass'],3), round(d['ms_per_step'],3))"; done; }
run base
export GLIA_B200_MINB=4; run minb4_e512
export GLIA_B200_ESLOTS=256; run minb4_e256
export GLIA_B200_TZ=16; run minb4_e256_tz16
unset GLIA_B200_MINB; run minb3_e256_tz16
