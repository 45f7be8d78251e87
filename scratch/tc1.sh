cd /root/repo
timeout 300 python scratch/tc1.py > gpurun_out/tc1.log 2>&1; echo "exit $?" >> gpurun_out/tc1.log
N=1048576 timeout 300 python scratch/tc1.py > gpurun_out/tc1_1m.log 2>&1; echo "exit $?" >> gpurun_out/tc1_1m.log
tail -30 gpurun_out/tc1.log gpurun_out/tc1_1m.log
