#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dichotomy or humanoid or edge_validation" 2>&1 | tail -n 25 > gpurun_out/t8_tests.log
cat gpurun_out/t8_tests.log
for m in 2; do
HPPB200_WF_TRACE=1 timeout 600 python tools/wf_run.py 1048576 0 2 2>&1 | grep -v "^\[wf\] round" | tail -n 6
HPPB200_WF_TRACE=1 timeout 600 python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4 2>&1 | grep -v "^\[wf\] round" | tail -n 6
done
