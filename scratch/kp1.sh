cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "knn or kprm" > gpurun_out/kp1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/kp1_pytest.log
tail -n 15 gpurun_out/kp1_pytest.log
timeout 900 python scratch/kp1.py > gpurun_out/kp1.log 2>&1; echo "exit $?" >> gpurun_out/kp1.log
cat gpurun_out/kp1.log | cut -c1-400
