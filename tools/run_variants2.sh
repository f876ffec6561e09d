#!/bin/bash
# variants in gpurun_variants/ + the in-tree library, on each workload in WLS (default "B")
mkdir -p gpurun_out
for wl in ${WLS:-B}; do
  B="python bench.py --workload $wl --steps 5 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
  for so in gpurun_variants/libshk_*.so HEAD; do
    if [ "$so" = HEAD ]; then v=head; unset SHK_LIB; else [ -f "$so" ] || continue; v=$(basename $so .so); v=${v#libshk_}; export SHK_LIB=$so; fi
    timeout 120 $B > gpurun_out/v2_${wl}_$v.json 2> gpurun_out/v2_${wl}_$v.err
    python - gpurun_out/v2_${wl}_$v.json <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    r=d["roofline"]
    print(sys.argv[1], round(d["ms_per_step"],3), "K1b", round(r["kernel_ms"],3), "K1a", round(r.get("k_minimizer_partition_ms") or 0,3), {k:round(v,3) for k,v in r["stage_ms"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
  done
done
