#!/bin/bash
# ncu passes of this round, each only after the same command exited 0 without ncu (B200_PROFILING.md).
# Usage: gpurun -- 'bash tools/prof_round.sh'   -> gpurun_out/r2_*.{csv,ncu-rep,log}
mkdir -p gpurun_out
set -x
# 1. launch list of the default bench command
python bench.py --steps 2 --warmup 3 > gpurun_out/r2_bench_plain.json 2> gpurun_out/r2_bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_bench_launches.csv \
    python bench.py --steps 2 --warmup 3 > gpurun_out/r2_bench_ncu.log 2>&1
# 2. full capture of the default kernel on the bench model and row count (retrained, 10 M rows)
export NCU_MODEL=retrained NCU_ROWS=10000000
python tools/ncu_target.py > gpurun_out/r2_target_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:hfe_mma -s 2 -c 1 -f -o gpurun_out/r2_mma_bench \
    python tools/ncu_target.py > gpurun_out/r2_target_ncu.log 2>&1
# 3. the streaming pass (no SVC member) on the same rows
export NCU_MODEL=nosvc
python tools/ncu_target.py > gpurun_out/r2_nosvc_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:hfe_stream -s 2 -c 1 -f -o gpurun_out/r2_stream_bench \
    python tools/ncu_target.py > gpurun_out/r2_nosvc_ncu.log 2>&1
# 4. the shared-memory forest walk on the stress forest, 2 M rows (bench leg size)
export NCU_MODEL=stress NCU_ROWS=2000000 NCU_LAUNCHES=3
python tools/ncu_target.py > gpurun_out/r2_stress_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:hfe_forest -s 1 -c 1 -f -o gpurun_out/r2_forest_bench \
    python tools/ncu_target.py > gpurun_out/r2_stress_ncu.log 2>&1
for f in gpurun_out/r2_*_ncu.log; do tail -n 2 $f; done
