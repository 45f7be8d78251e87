cd $GRAFT_REPO_ROOT
for v in "" scratch/variants/ch8/libhppb200.so scratch/variants/ch32/libhppb200.so scratch/variants/ch16r80/libhppb200.so; do
HPPB200_LIB=$v HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_p.txt 2>&1
echo "variant $v"; grep -E "parity|^n=|rounds=" gpurun_out/wf_p.txt | tail -3 | cut -c1-150
done
