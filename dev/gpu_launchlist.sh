#!/bin/bash
# ncu launch list of the step in its final state (encoders launched eagerly so every kernel is listed; dilation-18 conv folded)
mkdir -p gpurun_out/profiles_box
export CP2_BENCH_CUDA_PROFILER=1
STEP="python bench.py --steps 2 --warmup 3 --no-sweep --no-cpu-baseline --no-eager-baseline --no-graphs"
timeout 300 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
    --log-file gpurun_out/launches_step.csv $STEP > gpurun_out/ncu_step.log 2>&1
echo "== ncu launch list exit $?"
CP2_PROFILES_DST=gpurun_out/profiles_box python dev/summarise_profiles.py ${TAG:-r02h} > gpurun_out/summarise.log 2>&1; echo "== summarise exit $?"
head -16 gpurun_out/profiles_box/${TAG:-r02h}_launches_step.txt | cut -c1-140
