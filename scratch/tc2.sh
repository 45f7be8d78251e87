cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q -k "knn" > gpurun_out/tc2_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/tc2_pytest.log
tail -n 5 gpurun_out/tc2_pytest.log
timeout 200 python scratch/tc2.py > gpurun_out/tc2_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/tc2_launches.csv python scratch/tc2.py > gpurun_out/tc2_ncu1.log 2>&1
cat gpurun_out/tc2_plain.log
timeout 200 python scratch/tc2.py > gpurun_out/tc2_plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:knnTcFilter -s 2 -c 2 -o gpurun_out/tc2_filter python scratch/tc2.py > gpurun_out/tc2_ncu2.log 2>&1
ls -la gpurun_out
