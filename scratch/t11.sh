#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 600 python tools/tail_prof.py ur5_boxes8.json 4000 2>&1 | tail -n 2
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dichTailKernel -c 1 -o gpurun_out/r1n_tail python tools/tail_prof.py ur5_boxes8.json 4000 > gpurun_out/r1n_tail_ncu.log 2>&1
tail -n 3 gpurun_out/r1n_tail_ncu.log
