cd $GRAFT_REPO_ROOT
HPPB200_WF_TRACE=2 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_trace_prog.txt 2>&1
python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_plain_prog.txt 2>&1
HPPB200_WF_SLOTS=151552 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_plain_prog_E151k.txt 2>&1
HPPB200_WF_SLOTS=75776 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_plain_prog_E75k.txt 2>&1
tail -3 gpurun_out/wf_plain_prog*.txt
