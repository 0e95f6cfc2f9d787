#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t4_pytest.log
tail -8 gpurun_out/t4_pytest.log
run() { tag=$1; shift; timeout 600 python bench.py "$@" --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/t4_$tag.json 2> gpurun_out/t4_$tag.err; echo "bench $tag rc=$?"; }
run c4s --workload c4s
run c2 --workload c2
run c5 --workload c5
GLIA_BENCH_TZ=24 run c4s_tz24 --workload c4s
GLIA_BENCH_SLOTS=1024 GLIA_BENCH_TZ=24 run c4s_tz24_s1024 --workload c4s
GLIA_BENCH_TZ=24 run c2_tz24 --workload c2
GLIA_BENCH_TZ=31 run c2_tz31 --workload c2
python - <<'PY'
import json
for w in ("c4s","c2","c5","c4s_tz24","c4s_tz24_s1024","c2_tz24","c2_tz31"):
    try:
        d=json.loads(open("gpurun_out/t4_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "ms/step", round(d["ms_per_step"],3), {k:round(v,3) for k,v in d.get("stage_ms",{}).items()}, d.get("value"))
    except Exception as e: print(w,"fail",e)
PY
