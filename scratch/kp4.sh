cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "kprm or knn or config" > gpurun_out/kp4_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/kp4_pytest.log
tail -n 6 gpurun_out/kp4_pytest.log
timeout 1500 python scratch/kp4.py > gpurun_out/kp4.log 2>&1; echo "exit $?" >> gpurun_out/kp4.log
cat gpurun_out/kp4.log | cut -c1-700
