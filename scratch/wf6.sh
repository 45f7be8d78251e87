cd /root/repo
timeout 600 python tools/wf_run.py 131072 20000 > gpurun_out/wf6_a.log 2>&1
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf6_c.log 2>&1
HPPB200_WF_REFILL=16 HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf6_d.log 2>&1
HPPB200_WF_REFILL=32 HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf6_e.log 2>&1
HPPB200_WF_REFILL=1 HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf6_f.log 2>&1
for f in a c d e f; do echo == $f; grep -v "^n=\|wf\] round" gpurun_out/wf6_$f.log | head -n 4; grep "round 128" gpurun_out/wf6_$f.log | tail -n 1;  grep "^n=" gpurun_out/wf6_$f.log | tail -n 1 | cut -c1-330; done
