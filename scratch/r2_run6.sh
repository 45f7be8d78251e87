cd $GRAFT_REPO_ROOT
python scratch/r2_dbg.py 2>&1 | grep -E "^reverse|edge" | cut -c1-60
for v in "" scratch/variants/minb7/libhppb200.so scratch/variants/minb6/libhppb200.so; do
for k in 0 1.5 2.5 4; do
HPPB200_WF_KAPPA=$k HPPB200_LIB=$v HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_c2.txt 2>&1
echo "variant $v kappa $k"; grep -E "parity|^n=|rounds=" gpurun_out/wf_c2.txt | tail -2 | cut -c1-330
done; done
