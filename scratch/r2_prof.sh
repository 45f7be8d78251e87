cd $GRAFT_REPO_ROOT
python tools/prof_run.py configs > gpurun_out/prof_cfg_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"configPassKernel|gjkJobsKernel" -s 2 -c 2 -f -o gpurun_out/r2_cfg python tools/prof_run.py configs > gpurun_out/ncu_cfg.log 2>&1
tail -2 gpurun_out/ncu_cfg.log
python tools/prof_run.py prog 1048576 > gpurun_out/prof_prog_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/r2_prog_launches.csv python tools/prof_run.py prog 1048576 > gpurun_out/ncu_prog_list.log 2>&1
tail -2 gpurun_out/ncu_prog_list.log
ncu --set full --clock-control none --import-source on -k regex:"wfProgControl|wfNarrow|wfFk" -s 150 -c 3 -f -o gpurun_out/r2_wf python tools/prof_run.py prog 1048576 > gpurun_out/ncu_wf.log 2>&1
tail -2 gpurun_out/ncu_wf.log
