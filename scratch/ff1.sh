#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "freeflyer" 2>&1 | tail -n 25 > gpurun_out/ff1_tests.log
cat gpurun_out/ff1_tests.log
timeout 600 python bench.py --steps 3 --warmup 3 2> gpurun_out/ff1_bench.err | tee gpurun_out/ff1_bench.json | cut -c1-1500
