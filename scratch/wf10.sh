cd /root/repo
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf10_b.log 2>&1
grep "wf\]" gpurun_out/wf10_b.log | tail -n 20
