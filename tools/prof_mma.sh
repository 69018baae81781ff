#!/bin/bash
# full ncu capture of the tensor-core kernel on the bench model and row count (after the same command ran clean)
mkdir -p gpurun_out
export NCU_MODEL=retrained NCU_ROWS=10000000
timeout 120 python tools/ncu_target.py > gpurun_out/r2_target_plain.log 2>&1 || exit 1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:hfe_mma -s 2 -c 1 -f -o gpurun_out/r2_mma_bench \
    python tools/ncu_target.py > gpurun_out/r2_target_ncu.log 2>&1
tail -n 2 gpurun_out/r2_target_ncu.log
