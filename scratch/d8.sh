cd /root/repo
for tl in 0 2048; do HPPB200_WF_TAIL=$tl HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 2>&1 | grep "edges\]\|rounds=\|^n=" | tail -n 3 | cut -c1-150; done
