cd /root/repo
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 131072 20000 2 2>&1 | grep parity
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 2>&1 | grep "round 128\|rounds=\|^n=" | tail -n 3 | cut -c1-150
for m in 1 2; do python tools/edge_rate.py humanoid_obst6.json $m 1048576 0.4 2>&1 | grep method; done
python -m pytest tests -m gpu -x -q -k "wavefront or humanoid" 2>&1 | tail -n 3
