#!/bin/bash
# ncu evidence for bench.py (run under gpurun).  Plain run first, then the launch list, then one
# --set full capture per hot kernel.  Numbers printed under ncu are never bench values.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-same-box"
$CMD > gpurun_out/prof_plain.json 2> gpurun_out/prof_plain.err || { echo "plain run failed"; tail gpurun_out/prof_plain.err; exit 1; }
tail -c 400 gpurun_out/prof_plain.json; echo
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$? lines=$(wc -l < gpurun_out/launches.csv)"
ncu --set full --clock-control none --import-source on -k regex:ppm_attend -s 20 -c 1 -f -o gpurun_out/prof_attend $CMD > gpurun_out/ncu_attend.log 2>&1
echo "attend rc=$?"
ncu --set full --clock-control none --import-source on -k regex:corr_lookup -s 20 -c 1 -f -o gpurun_out/prof_lookup $CMD > gpurun_out/ncu_lookup.log 2>&1
echo "lookup rc=$?"
ncu --set full --clock-control none --import-source on -k regex:corr_build -s 2 -c 1 -f -o gpurun_out/prof_build $CMD > gpurun_out/ncu_build.log 2>&1
echo "build rc=$?"
ncu --set full --clock-control none --import-source on -k regex:pack_key_kernel -s 30 -c 1 -f -o gpurun_out/prof_pack $CMD > gpurun_out/ncu_pack.log 2>&1
echo "pack rc=$?"
ls -la gpurun_out/*.ncu-rep
