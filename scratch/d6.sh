cd /root/repo
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 131072 20000 2 > gpurun_out/d6_a.log 2>&1
grep -v "^n=" gpurun_out/d6_a.log | head -n 6; grep "^n=" gpurun_out/d6_a.log | tail -n 1 | cut -c1-330
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 > gpurun_out/d6_d.log 2>&1
grep "round 128\|rounds=" gpurun_out/d6_d.log | tail -n 2; grep "^n=" gpurun_out/d6_d.log | tail -n 1 | cut -c1-330
for sl in 303104 75776; do HPPB200_WF_SLOTS=$sl timeout 300 python tools/wf_run.py 1048576 0 2 2>&1 | grep "^n=" | tail -n 1 | cut -c1-120; done
