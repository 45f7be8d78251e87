cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu -k "wavefront_engine or humanoid or freeflyer or birrt or pair_table or chains" 2>&1 | tail -3
HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_p.txt 2>&1
grep -E "parity|^n=|rounds=" gpurun_out/wf_p.txt | tail -4 | cut -c1-250
