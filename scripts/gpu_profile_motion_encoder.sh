#!/bin/bash
# ncu evidence for K7 (run under gpurun): plain run first, then a launch list of one encoder call sequence and a
# --set full capture of K6b (channels-last store) and the three 3x3 shifted GEMMs.
mkdir -p gpurun_out
CMD="python scripts/bench_motion_encoder.py 20 184 320 --kernel-only"
python scripts/bench_motion_encoder.py 20 184 320 > gpurun_out/prof_menc_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/prof_menc_plain.log; exit 1; }
tail -1 gpurun_out/prof_menc_plain.log
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_menc.csv $CMD > gpurun_out/ncu_menc_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < gpurun_out/launches_menc.csv)"
ncu --set full --clock-control none --import-source on -k regex:'gru_gemm|motion_front_b' -s 4 -c 4 -f -o gpurun_out/prof_menc $CMD > gpurun_out/ncu_menc.log 2>&1
echo "menc rc=$?"
ls -la gpurun_out/*.ncu-rep
