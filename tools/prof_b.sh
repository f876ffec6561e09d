#!/bin/bash
# ncu --set full captures of the two counting kernels on config B (source-level counters), after a plain run exited 0
mkdir -p gpurun_out
A="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
timeout 120 python bench.py $A > gpurun_out/pb_plain.json 2> gpurun_out/pb_plain.err || exit 1
timeout 300 ncu --set full --import-source on --clock-control none -k regex:"k_count_buckets|k_minimizer_partition" --launch-skip 9 --launch-count 3 -o gpurun_out/prof_b -f python bench.py $A > gpurun_out/pb_ncu.log 2>&1
ls -la gpurun_out/prof_b.ncu-rep
