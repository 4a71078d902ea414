#!/bin/bash
# Run the GPU parity suites in separate processes (a device trap in one suite must not hide the others).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.used --format=csv > gpurun_out/gpu.txt 2>&1
run() { name=$1; shift; timeout 900 "$@" > gpurun_out/$name.log 2>&1; echo "== $name exit $?"; tail -n 25 gpurun_out/$name.log; }
run smoke python __graft_entry__.py --smoke
run corr python -m pytest tests/test_corr_gpu.py -q -m gpu -p no:cacheprovider
run ppm_tmem python -m pytest tests/test_ppm_read_gpu.py -q -m gpu -p no:cacheprovider -k "not False"
run ppm_smem python -m pytest tests/test_ppm_read_gpu.py -q -m gpu -p no:cacheprovider -k "False"
