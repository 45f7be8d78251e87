cd $GRAFT_REPO_ROOT
python scratch/r2_dbg.py 2>&1 | grep -E "^reverse|edge" | cut -c1-60
for v in "" scratch/variants/minb6/libhppb200.so; do
HPPB200_LIB=$v HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_c2.txt 2>&1
echo "variant $v"; grep -E "parity|^n=|rounds=" gpurun_out/wf_c2.txt | tail -4
done
ncu --set full --import-source on --clock-control none -k regex:wfProgControl --launch-skip 40 --launch-count 1 -o gpurun_out/r2_ctl4 -f python tools/prof_run.py prog 1048576 > gpurun_out/ncu_ctl4.log 2>&1
tail -2 gpurun_out/ncu_ctl4.log
