#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/t3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t3_pytest.log
tail -8 gpurun_out/t3_pytest.log
for w in c4s c2 c4; do
  timeout 600 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/t3_bench_$w.json 2> gpurun_out/t3_bench_$w.err; echo "bench $w rc=$?"
done
python - <<'PY'
import json
for w in ("c4s","c2","c4"):
    try:
        d=json.loads(open("gpurun_out/t3_bench_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "ms/step", round(d["ms_per_step"],3), {k:round(v,3) for k,v in d["stage_ms"].items()})
    except Exception as e: print(w,"fail",e)
PY
CMD="python bench.py --workload c4s --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/t3_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:voxel_pass_fast -s 2 -c 1 -o gpurun_out/t3_voxel $CMD > gpurun_out/t3_ncu.log 2>&1
echo "ncu rc=$?"
