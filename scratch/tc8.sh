cd /root/repo
N=1048576 timeout 300 python scratch/tc1.py > gpurun_out/tc8_1m.log 2>&1; echo "exit $?" >> gpurun_out/tc8_1m.log
tail -n 14 gpurun_out/tc8_1m.log
timeout 900 python -m pytest tests -m gpu -x -q -k "knn" > gpurun_out/tc8_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/tc8_pytest.log
tail -n 15 gpurun_out/tc8_pytest.log
timeout 200 python scratch/tc2.py > gpurun_out/tc8_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tc8_launches.csv python scratch/tc2.py > gpurun_out/tc8_ncu1.log 2>&1
cat gpurun_out/tc8_plain.log
