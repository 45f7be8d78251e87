cd $GRAFT_REPO_ROOT
for v in "" scratch/variants/nminb4/libhppb200.so scratch/variants/nminb2/libhppb200.so; do
HPPB200_LIB=$v HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_p.txt 2>&1
echo "variant $v"; grep -E "^n=|rounds=" gpurun_out/wf_p.txt | tail -2 | cut -c1-150
done
