#!/bin/bash
cd $GRAFT_REPO_ROOT
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 > gpurun_out/r1n_pytest.log; cat gpurun_out/r1n_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1 | cut -c1-200
timeout 900 python bench.py > gpurun_out/r1n_bench.json 2> gpurun_out/r1n_bench.err; tail -n 2 gpurun_out/r1n_bench.err
(HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py ur5_boxes8.json 20000; HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py humanoid_obst6.json 20000 0.4; HPPB200_WF_TRACE=1 timeout 600 python tools/tail_latency.py humanoid_ff_obst6.json 20000 0.4) 2>&1 | grep -v "^\[edges\]" | awk '!seen[$0]++' > gpurun_out/r1n_tail_latency.txt
tail -n 4 gpurun_out/r1n_tail_latency.txt | cut -c1-250
(timeout 600 python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4; timeout 600 python tools/edge_rate.py humanoid_ff_obst6.json 2 1048576 0.4; timeout 600 python tools/edge_rate.py humanoid_ff_obst6.json 1 1048576 0.4; timeout 600 python tools/edge_rate.py humanoid_ff_obst6.json 0 262144 0.4) 2>&1 | grep method > gpurun_out/r1n_edge_rates_humanoid.txt
cat gpurun_out/r1n_edge_rates_humanoid.txt | cut -c1-250
timeout 900 ncu --set full --clock-control none --import-source on -k regex:dichTailKernel -c 1 -f -o gpurun_out/r1n_tail python tools/tail_prof.py ur5_boxes8.json 4000 > gpurun_out/r1n_tail_ncu.log 2>&1; tail -n 1 gpurun_out/r1n_tail_ncu.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1n_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-edges --no-cpu > gpurun_out/r1n_ncu_bench.log 2>&1; wc -l gpurun_out/r1n_launches_bench_steps2.csv
