cd /root/repo
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 2>&1 | grep "edges\]\|rounds=\|^n=" | tail -n 3 | cut -c1-150
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 131072 20000 2 2>&1 | grep "parity"
