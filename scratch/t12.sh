#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dichotomy or humanoid" 2>&1 | tail -n 5
HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py ur5_boxes8.json 20000 2>&1 | grep -v "^\[edges\]" | tail -n 5
HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py humanoid_obst6.json 20000 0.4 2>&1 | grep -v "^\[edges\]" | tail -n 5
HPPB200_WF_TRACE=1 timeout 600 python tools/wf_run.py 1048576 0 2 2>&1 | grep -v "^\[wf\] round" | tail -n 4
HPPB200_WF_TRACE=1 timeout 600 python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4 2>&1 | grep -v "^\[wf\] round" | tail -n 4
