cd /root/repo
timeout 600 python tools/wf_run.py 131072 20000 > gpurun_out/wf8_a.log 2>&1
HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf8_m4.log 2>&1
for mb in 3 2; do
  rm -f hpp-core_b200/csrc/build/kernels_wavefront.o
  make -C hpp-core_b200/csrc EXTRA=-DWF_NMINB=$mb > /dev/null 2>&1
  HPPB200_WF_TRACE=1 timeout 300 python tools/wf_run.py 1048576 0 > gpurun_out/wf8_m$mb.log 2>&1
done
for f in a m4 m3 m2; do echo == $f; grep -v "^n=\|wf\] round" gpurun_out/wf8_$f.log | head -n 3; grep "round 128" gpurun_out/wf8_$f.log | tail -n 1;  grep "^n=" gpurun_out/wf8_$f.log | tail -n 1 | cut -c1-330; done
