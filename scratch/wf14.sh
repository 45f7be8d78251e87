cd /root/repo
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf14_a.log 2>&1
rm -f hpp-core_b200/csrc/build/kernels_wavefront.o
make -C hpp-core_b200/csrc EXTRA=-DWF_WARP_ATOMICS > /dev/null 2>&1
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf14_b.log 2>&1
for f in a b; do echo == $f; grep "round 128" gpurun_out/wf14_$f.log | tail -n 1;  grep "^n=" gpurun_out/wf14_$f.log | tail -n 1 | cut -c1-100; done
timeout 800 ncu --set full --clock-control none --import-source on -k regex:wfControl --launch-skip 100 --launch-count 1 -o gpurun_out/wf14c -f python tools/wf_run.py 1048576 0 > gpurun_out/wf14.log 2>&1
