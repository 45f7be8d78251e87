cd $GRAFT_REPO_ROOT
for d in 0 1; do
HPPB200_WF_DIRECT=$d HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_p.txt 2>&1
echo direct $d; grep -E "parity|^n=|rounds=|continue" gpurun_out/wf_p.txt | tail -3 | cut -c1-250
done
