#!/bin/bash
# ncu --set full captures of K1 (pyramid build) and K2 (lookup) at the 720p 1/4-res shape, plus the
# lookup's DRAM traffic under the three L2 fetch granularities (run under gpurun).
mkdir -p gpurun_out
CMD="python scripts/bench_corr.py"
$CMD > gpurun_out/corr_plain.log 2>&1 || { echo "plain run failed"; tail gpurun_out/corr_plain.log; exit 1; }
cat gpurun_out/corr_plain.log
ncu --set full --clock-control none --import-source on -k regex:corr_build -s 2 -c 1 -f -o gpurun_out/prof_build $CMD > gpurun_out/ncu_build.log 2>&1
echo "build rc=$?"
ncu --set full --clock-control none --import-source on -k regex:corr_lookup -s 2 -c 1 -f -o gpurun_out/prof_lookup $CMD > gpurun_out/ncu_lookup.log 2>&1
echo "lookup rc=$?"
for gran in 32 64 128; do
  PPM_L2_FETCH=$gran ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:corr_lookup -s 2 -c 1 --csv --log-file gpurun_out/lookup_gran$gran.csv $CMD > /dev/null 2>&1
  echo "gran $gran:"; grep -h "dram__bytes\|gpu__time" gpurun_out/lookup_gran$gran.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
  PPM_L2_FETCH=$gran $CMD | grep K2
done
