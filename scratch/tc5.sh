cd /root/repo
N=1048576 timeout 300 python scratch/tc1.py > gpurun_out/tc5_1m.log 2>&1; echo "exit $?" >> gpurun_out/tc5_1m.log
cat gpurun_out/tc5_1m.log | tail -n 20
timeout 900 python -m pytest tests -m gpu -x -q -k "knn" > gpurun_out/tc5_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/tc5_pytest.log
tail -n 15 gpurun_out/tc5_pytest.log
timeout 200 python scratch/tc2.py > gpurun_out/tc5_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tc5_launches.csv python scratch/tc2.py > gpurun_out/tc5_ncu1.log 2>&1
cat gpurun_out/tc5_plain.log
timeout 200 python scratch/tc2.py > gpurun_out/tc5_plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:knnTcFilter -s 3 -c 3 -o gpurun_out/tc5_filter python scratch/tc2.py > gpurun_out/tc5_ncu2.log 2>&1
ls gpurun_out
