cd /root/repo
for tl in 0 256; do HPPB200_WF_TAIL=$tl HPPB200_WF_TRACE=1 python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4 2>&1 | grep "method\|edges\]\|rounds=" | tail -n 3 | cut -c1-200; done
