#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "dichotomy or humanoid" 2>&1 | tail -n 5
for sm in 1 0; do
echo "== TAIL_SMEM=$sm"
HPPB200_TAIL_SMEM=$sm HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py ur5_boxes8.json 20000 2>&1 | grep -v "^\[edges\]" | tail -n 4
HPPB200_TAIL_SMEM=$sm HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py humanoid_obst6.json 20000 0.4 2>&1 | grep -v "^\[edges\]" | tail -n 4
done
