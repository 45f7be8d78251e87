#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 600 python tools/tail_latency.py ur5_boxes8.json 20000 2>&1 | tail -n 8
timeout 600 python tools/tail_latency.py humanoid_obst6.json 20000 0.4 2>&1 | tail -n 8
timeout 600 python tools/tail_latency.py humanoid_ff_obst6.json 20000 0.4 2>&1 | tail -n 8
