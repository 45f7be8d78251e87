cd /root/repo
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 131072 20000 2 > gpurun_out/d1_a.log 2>&1
grep -v "^n=" gpurun_out/d1_a.log | head -n 20; grep "^n=" gpurun_out/d1_a.log | tail -n 1 | cut -c1-330
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 > gpurun_out/d1_b.log 2>&1
grep "round 64 \|round 128\|rounds=" gpurun_out/d1_b.log | tail -n 3; grep "^n=" gpurun_out/d1_b.log | tail -n 1 | cut -c1-330
