cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "knn or kprm or birrt" > gpurun_out/kp2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/kp2_pytest.log
tail -n 15 gpurun_out/kp2_pytest.log
timeout 1200 python scratch/kp1.py > gpurun_out/kp2.log 2>&1; echo "exit $?" >> gpurun_out/kp2.log
grep -v "^k " gpurun_out/kp2.log | cut -c1-400
