#!/bin/bash
# ncu evidence for K5 (run under gpurun): plain run first, then a launch list and one --set full capture of the first
# four GEMM launches of a call (conv z1|r1 first layer N=256 K=7680; z second layer; r second layer; q).
mkdir -p gpurun_out
CMD="python scripts/bench_gru.py 20 184 320"
$CMD > gpurun_out/prof_gru_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/prof_gru_plain.log; exit 1; }
tail -1 gpurun_out/prof_gru_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gru_ --csv --log-file gpurun_out/launches_gru.csv $CMD > gpurun_out/ncu_gru_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < gpurun_out/launches_gru.csv)"
ncu --set full --clock-control none --import-source on -k regex:gru_gemm -s 8 -c 8 -f -o gpurun_out/prof_gru $CMD > gpurun_out/ncu_gru.log 2>&1
echo "gru rc=$?"
ls -la gpurun_out/*.ncu-rep
