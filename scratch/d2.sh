cd /root/repo
rm -f hpp-core_b200/csrc/build/kernels_wavefront.o
make -C hpp-core_b200/csrc EXTRA=-DWF_DEBUG_OVERFLOW > /dev/null 2>&1
HPPB200_WF_MIN=1 timeout 600 python tools/wf_run.py 4096 0 2 > gpurun_out/d2_a.log 2>&1
grep "overflow" gpurun_out/d2_a.log | head -n 30; grep -c "pair overflow" gpurun_out/d2_a.log; grep -c "subset overflow" gpurun_out/d2_a.log
