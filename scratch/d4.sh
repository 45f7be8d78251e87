cd /root/repo
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 > gpurun_out/d4_d.log 2>&1
grep "round 128\|rounds=" gpurun_out/d4_d.log | tail -n 2; grep "^n=" gpurun_out/d4_d.log | tail -n 1 | cut -c1-330
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 1 > gpurun_out/d4_p.log 2>&1
grep "round 128\|rounds=" gpurun_out/d4_p.log | tail -n 2; grep "^n=" gpurun_out/d4_p.log | tail -n 1 | cut -c1-330
