#!/bin/bash
# On the GPU box (under gpurun): the round's bench lines and ncu captures into gpurun_out/ (copied to profiles/ afterwards
# with tools/summarize_launches.py / tools/summarize_ncu.py).  usage: bash tools/refresh_profiles.sh [tag]
tag=${1:-r02}
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err || echo "bench failed"
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${tag}_reference_arm.json 2> gpurun_out/${tag}_reference_arm.err || echo "reference arm failed"
# launch list of a timed region (the command has just exited 0 without ncu)
timeout 300 python bench.py --steps 2 --warmup 3 --repeats 1 --no-train --no-image-in --no-cpu-baseline --sat-batch 0 > /dev/null 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${tag}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 3 --repeats 1 --no-train --no-image-in --no-cpu-baseline --sat-batch 0 > gpurun_out/${tag}_ncu_launches.log 2>&1
# one eager step, full sections
timeout 120 python tools/run_step_once.py > /dev/null 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${tag}_full python tools/run_step_once.py > gpurun_out/${tag}_ncu_full.log 2>&1
ls -la gpurun_out | tail -12
cut -c 1-400 gpurun_out/${tag}_bench.json
cut -c 1-300 gpurun_out/${tag}_reference_arm.json
