cd /root/repo
for a in 4 1e9; do
HPPB200_ENGINE_ALPHA=$a HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 2 > gpurun_out/d3_$a.log 2>&1
echo alpha $a; grep "round 128\|rounds=" gpurun_out/d3_$a.log | tail -n 2; grep "^n=" gpurun_out/d3_$a.log | tail -n 1 | cut -c1-330
done
