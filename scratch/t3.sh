cd /root/repo
python -m pytest tests -m gpu -x -q -k humanoid 2>&1 | tail -n 30
