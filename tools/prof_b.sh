#!/bin/bash
# ncu --set full captures of the two counting kernels on config B (source-level counters), after a plain run exited 0
mkdir -p gpurun_out
A="--steps 1 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
W=${WL:-B}
timeout 120 python bench.py --workload $W $A > gpurun_out/pb_plain.json 2> gpurun_out/pb_plain.err || exit 1
# K1b: two launches per step (the plain and the spill-aware twin; one of them leaves at once): the fourth step's pair
timeout 300 ncu --set full --import-source on --clock-control none -k regex:"k_count_buckets" --launch-skip 6 --launch-count 2 -o gpurun_out/prof_k1b -f python bench.py --workload $W $A > gpurun_out/pb_ncu1.log 2>&1
# K1a: the last launches of the run are the timed step's (a host load cuts the reads slice by slice: many short launches before)
timeout 300 ncu --set full --import-source on --clock-control none -k regex:"k_minimizer_partition" --launch-skip ${K1A_SKIP:-3} --launch-count 1 -o gpurun_out/prof_k1a -f python bench.py --workload $W $A > gpurun_out/pb_ncu2.log 2>&1
ls -la gpurun_out/prof_k1b.ncu-rep gpurun_out/prof_k1a.ncu-rep
