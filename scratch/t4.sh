cd /root/repo
for m in 1 2 0; do python tools/edge_rate.py humanoid_obst6.json $m 262144 0.4 2>&1 | grep method; done
for m in 1 2; do python tools/edge_rate.py humanoid_obst6.json $m 262144 2>&1 | grep method; done
python tools/edge_rate.py ur5_boxes8.json 2 262144 2>&1 | grep method
python tools/edge_rate.py ur5_boxes8.json 1 16384 2>&1 | grep method
HPPB200_WF_MIN=1 python tools/edge_rate.py ur5_boxes8.json 1 16384 2>&1 | grep method
