cd /root/repo
timeout 600 python tools/wf_run.py 131072 20000 > gpurun_out/wf3_a.log 2>&1
timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf3_c.log 2>&1
HPPB200_WF_SLOTS=151552 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf3_d.log 2>&1
HPPB200_WF_SLOTS=606208 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf3_e.log 2>&1
HPPB200_ENGINE_ALPHA=1.2 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf3_f.log 2>&1
HPPB200_ENGINE_ALPHA=4 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf3_g.log 2>&1
for f in a c d e f g; do echo == $f; grep -v "^n=" gpurun_out/wf3_$f.log | head -n 5; grep "^n=" gpurun_out/wf3_$f.log | tail -n 1 | cut -c1-330; done
