cd /root/repo
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 131072 20000 2 > gpurun_out/d5_a.log 2>&1
grep -v "^n=" gpurun_out/d5_a.log | head -n 6; grep "^n=" gpurun_out/d5_a.log | tail -n 1 | cut -c1-330
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 > gpurun_out/d5_d.log 2>&1
grep "round 128\|rounds=" gpurun_out/d5_d.log | tail -n 2; grep "^n=" gpurun_out/d5_d.log | tail -n 1 | cut -c1-330
timeout 300 python tools/wf_run.py 1048576 20000 1 > gpurun_out/d5_p.log 2>&1
grep -v "^n=" gpurun_out/d5_p.log | head -n 4; grep "^n=" gpurun_out/d5_p.log | tail -n 1 | cut -c1-330
for tl in 512 8192; do HPPB200_WF_TAIL=$tl timeout 300 python tools/wf_run.py 1048576 0 2 2>&1 | grep "^n=" | tail -n 1 | cut -c1-120; done
for tl in 0 512 8192; do HPPB200_WF_TAIL=$tl timeout 300 python tools/wf_run.py 1048576 0 1 2>&1 | grep "^n=" | tail -n 1 | cut -c1-120; done
