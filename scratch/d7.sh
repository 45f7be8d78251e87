cd /root/repo
timeout 800 ncu --set full --clock-control none --import-source on -k regex:wfDichControl --launch-skip 100 --launch-count 1 -o gpurun_out/d7c -f python tools/wf_run.py 1048576 0 2 > gpurun_out/d7.log 2>&1
ls -la gpurun_out/d7c*
