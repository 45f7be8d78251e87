cd /root/repo
timeout 200 python scratch/tc2.py > gpurun_out/tc7_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:knnTcRerank -s 1 -c 3 -o gpurun_out/tc7_rerank python scratch/tc2.py > gpurun_out/tc7_ncu.log 2>&1
ls gpurun_out | grep tc7
