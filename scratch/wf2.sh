cd /root/repo
timeout 800 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/wf2_launches.csv python tools/wf_run.py 131072 0 > gpurun_out/wf2.log 2>&1
tail -n 3 gpurun_out/wf2.log
