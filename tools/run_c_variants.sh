#!/bin/bash
# build-knob variants of K1b (gpurun_variants/libshk_*.so) and table loads (LOADS="0.45 0.6") against the in-tree
# library, on workload WL (default C)
mkdir -p gpurun_out
B="python bench.py --workload ${WL:-C} --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
for so in gpurun_variants/libshk_*.so; do
  [ -f "$so" ] || continue
  v=$(basename $so .so); v=${v#libshk_}
  SHK_LIB=$so timeout 120 $B > gpurun_out/cv_$v.json 2> gpurun_out/cv_$v.err
done
for ld in ${LOADS:-}; do
  SHK_LOAD=$ld timeout 120 $B > gpurun_out/cv_load$ld.json 2> gpurun_out/cv_load$ld.err
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
