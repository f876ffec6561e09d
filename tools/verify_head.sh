#!/bin/bash
# one GPU call: the GPU test suite, the default bench line, the ncu launch list of the headline step, build-knob variants
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/vh_gpu.txt
nproc >> gpurun_out/vh_gpu.txt
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/vh_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/vh_pytest.log
timeout 600 python bench.py > gpurun_out/vh_bench_B_n1.json 2> gpurun_out/vh_bench_B_n1.err; echo "bench rc=$?"
A="--steps 2 --warmup 3 --no-e2e --no-cpu-baseline --no-cold --no-extras"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/vh_launches_B.csv python bench.py $A > gpurun_out/vh_ncu.log 2>&1
WL=B LOADS="${LOADS:-0.45 0.65}" tools/run_c_variants.sh > gpurun_out/vh_variants.txt 2>&1
tail -3 gpurun_out/vh_pytest.log; cat gpurun_out/vh_variants.txt
