cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/kp3_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/kp3_pytest.log
tail -n 6 gpurun_out/kp3_pytest.log
timeout 1200 python scratch/kp3.py > gpurun_out/kp3.log 2>&1; echo "exit $?" >> gpurun_out/kp3.log
cat gpurun_out/kp3.log | cut -c1-500
