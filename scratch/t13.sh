#!/bin/bash
cd $GRAFT_REPO_ROOT
for tm in 1000000000 0; do
echo "== TAIL_MAX=$tm"
HPPB200_TAIL_MAX=$tm timeout 600 python tools/edge_rate.py ur5_boxes8.json 2 200000 2>&1 | tail -n 1
HPPB200_TAIL_MAX=$tm timeout 600 python tools/edge_rate.py humanoid_obst6.json 2 200000 0.4 2>&1 | tail -n 1
HPPB200_TAIL_MAX=$tm timeout 600 python tools/edge_rate.py ur5_boxes8.json 2 10000 2>&1 | tail -n 1
done
echo "== wavefront at 200k"
HPPB200_WF_MIN=1 HPPB200_TAIL_MAX=0 timeout 600 python tools/edge_rate.py ur5_boxes8.json 2 200000 2>&1 | tail -n 1
HPPB200_WF_MIN=1 HPPB200_TAIL_MAX=0 timeout 600 python tools/edge_rate.py humanoid_obst6.json 2 200000 0.4 2>&1 | tail -n 1
HPPB200_WF_MIN=1 HPPB200_TAIL_MAX=0 timeout 600 python tools/edge_rate.py ur5_boxes8.json 2 50000 2>&1 | tail -n 1
