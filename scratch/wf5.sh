cd /root/repo
timeout 800 ncu --set full --clock-control none --import-source on -k regex:wf --launch-skip 201 --launch-count 2 -o gpurun_out/wf5 -f python tools/wf_run.py 1048576 0 > gpurun_out/wf5.log 2>&1
tail -n 3 gpurun_out/wf5.log | cut -c1-200
ls -la gpurun_out/wf5*
