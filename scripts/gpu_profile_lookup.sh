#!/bin/bash
# ncu --set full capture of K2b (lookup from level 0) at the 720p 1/4-res shape (run under gpurun).
mkdir -p gpurun_out
CMD="python scripts/bench_corr.py"
$CMD > gpurun_out/corr_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/corr_plain.log; exit 1; }
cat gpurun_out/corr_plain.log
ncu --set full --clock-control none --import-source on -k regex:corr_lookup_l0 -s 2 -c 1 -f -o gpurun_out/prof_lookup_l0 $CMD > gpurun_out/ncu_lookup_l0.log 2>&1
echo "lookup_l0 rc=$?"
