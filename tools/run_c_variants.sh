#!/bin/bash
# config C (k = 63): build-knob variants of K1b (gpurun_variants/libshk_*.so) against the in-tree library
mkdir -p gpurun_out
B="python bench.py --workload C --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
for so in gpurun_variants/libshk_*.so; do
  v=$(basename $so .so); v=${v#libshk_}
  SHK_LIB=$so timeout 120 $B > gpurun_out/cv_$v.json 2> gpurun_out/cv_$v.err
done
timeout 120 $B > gpurun_out/cv_head.json 2> gpurun_out/cv_head.err
for f in gpurun_out/cv_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], round(d["ms_per_step"],3), round(d["roofline"]["kernel_ms"],3), d["roofline"]["stage_ms"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
