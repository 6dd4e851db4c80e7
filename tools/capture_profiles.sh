#!/bin/bash
# GPU box: the round's evidence in one call -- bench line, ncu launch list of the same command, one --set full capture of
# the loop kernel (each ncu pass only after its command has exited 0 without ncu).  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
python bench.py > gpurun_out/r2_bench_line.json 2> gpurun_out/r2_bench.err || exit 1
python bench.py --steps 2 --warmup 3 --no-extras > /dev/null 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r2_ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:sample_loop_kernel -s 3 -c 1 -o gpurun_out/r2_loop \
    python bench.py --steps 1 --warmup 3 --no-extras > gpurun_out/r2_ncu_full.log 2>&1
python tools/bench_variants.py c3 c4 c5 > gpurun_out/r2_variants.jsonl 2>&1
