#!/bin/bash
# scratch: first GPU check of the round-2 voxel kernel
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/t1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t1_pytest.log
tail -5 gpurun_out/t1_pytest.log
for w in c4s c2; do
  timeout 300 python bench.py --workload $w --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/t1_bench_$w.json 2> gpurun_out/t1_bench_$w.err; echo "bench $w rc=$?"
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/t1_bench_c4.json 2> gpurun_out/t1_bench_c4.err; echo "bench c4 rc=$?"
python - <<'PY'
import json
for w in ("c4s","c2","c4"):
    try:
        d=json.loads(open("gpurun_out/t1_bench_%s.json"%w).read().strip().splitlines()[-1])
        print(w, "ms/step", round(d["ms_per_step"],3), d["stage_ms"], d["config"].get("forest"))
    except Exception as e: print(w,"fail",e)
PY
