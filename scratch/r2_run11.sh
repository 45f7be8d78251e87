cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu -k "wavefront or humanoid or freeflyer or full_size or birrt or dichotomy or kprm" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -3 gpurun_out/r2_pytest_wf.txt
HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_p.txt 2>&1
grep -E "parity|^n=|rounds=|continue" gpurun_out/wf_p.txt | tail -5 | cut -c1-250
python tools/wf_run.py 1048576 0 2 > gpurun_out/wf_d.txt 2>&1
grep -E "parity|^n=|rounds=|continue" gpurun_out/wf_d.txt | tail -1 | cut -c1-250
