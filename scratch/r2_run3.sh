cd $GRAFT_REPO_ROOT
python scratch/r2_dbg.py 2>&1 | grep -E "^reverse" 
python -m pytest tests -x -q -m gpu -k "wavefront_engine or humanoid or freeflyer or full_size or birrt" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -3 gpurun_out/r2_pytest_wf.txt
ncu --set full --import-source on --clock-control none -k regex:wfProgControl --launch-skip 40 --launch-count 1 -o gpurun_out/r2_ctl2 -f python tools/prof_run.py prog 400000 > gpurun_out/ncu_ctl2.log 2>&1
tail -2 gpurun_out/ncu_ctl2.log
