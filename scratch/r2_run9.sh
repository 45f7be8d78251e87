cd $GRAFT_REPO_ROOT
python scratch/r2_small.py 2>&1 | tail -12
python -m pytest tests -x -q -m gpu -k "wavefront or humanoid or freeflyer or full_size or birrt or dichotomy" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -3 gpurun_out/r2_pytest_wf.txt
for h in 4096 0 1024 16384; do
HPPB200_WF_HAND=$h HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_h$h.txt 2>&1
echo "prog hand $h"; grep -E "parity|^n=|rounds=|continue|tail" gpurun_out/wf_h$h.txt | tail -5 | cut -c1-330
done
