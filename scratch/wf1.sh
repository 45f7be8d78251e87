set -x
cd /root/repo
timeout 600 python tools/wf_run.py 131072 20000 > gpurun_out/wf1_a.log 2>&1
HPPB200_ENGINE=1 timeout 300 python tools/wf_run.py 131072 0 > gpurun_out/wf1_b.log 2>&1
timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf1_c.log 2>&1
HPPB200_ENGINE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf1_d.log 2>&1
tail -5 gpurun_out/wf1_*.log
