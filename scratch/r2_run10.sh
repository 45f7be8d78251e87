cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest_all.txt 2>&1; tail -3 gpurun_out/r2_pytest_all.txt
for h in 2048 4096 8192; do
HPPB200_WF_HAND=$h HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_h$h.txt 2>&1
echo "prog hand $h"; grep -E "parity|^n=|rounds=|continue" gpurun_out/wf_h$h.txt | tail -5 | cut -c1-200
done
