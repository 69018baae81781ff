#!/bin/bash
# One GPU-box pass: the GPU test tier, then the default bench line.  Usage: gpurun -- 'bash tools/gpu_round.sh TAG'
tag=${1:-cur}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_gpu_tests.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/${tag}_gpu_tests.log
tail -5 gpurun_out/${tag}_gpu_tests.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${tag}_smoke.log
timeout 900 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err
echo "bench exit $?"
cat gpurun_out/${tag}_bench.json
tail -5 gpurun_out/${tag}_bench.err
