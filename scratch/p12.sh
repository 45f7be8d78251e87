cd /root/repo
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
export HPPB200_KNN_TC_PREFILTER=1
timeout 900 python -m pytest tests -m gpu -x -q -k "knn or kprm" 2>&1 | tail -n 3
N=1048576 timeout 300 python scratch/tc1.py 2>&1 | tail -n 5
timeout 200 python scratch/tc2.py > gpurun_out/r1u_tc2_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1u_knnTc_launches.csv python scratch/tc2.py > gpurun_out/r1u_ncu1.log 2>&1
grep -c Rerank gpurun_out/r1u_knnTc_launches.csv
