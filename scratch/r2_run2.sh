cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu -k "wavefront_engine or humanoid or freeflyer or full_size or birrt" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -5 gpurun_out/r2_pytest_wf.txt
HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_c2.txt 2>&1
grep -E "parity|^n=|rounds=" gpurun_out/wf_c2.txt | tail -8
HPPB200_WF_CONTROL=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_c1.txt 2>&1
grep -E "^n=" gpurun_out/wf_c1.txt | tail -1
