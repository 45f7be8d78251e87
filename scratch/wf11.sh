cd /root/repo
timeout 600 python tools/wf_run.py 131072 20000 > gpurun_out/wf16_a.log 2>&1
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf16_b.log 2>&1
for f in a b; do echo == $f; grep -v "^n=\|wf\] round" gpurun_out/wf16_$f.log | head -n 3; grep "round 128" gpurun_out/wf16_$f.log | tail -n 1;  grep "^n=" gpurun_out/wf16_$f.log | tail -n 1 | cut -c1-330; done
