cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "knn or kprm or birrt" 2>&1 | tail -n 3
N=1048576 timeout 300 python scratch/tc1.py 2>&1 | tail -n 6
timeout 300 python scratch/kp1.py 2>&1 | grep "^k " | cut -c1-200
