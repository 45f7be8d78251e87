cd $GRAFT_REPO_ROOT
for e in 0 227328 151552 113664 75776; do
HPPB200_WF_SLOTS=$e HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_p.txt 2>&1
echo slots $e; grep -E "^n=|rounds=" gpurun_out/wf_p.txt | tail -2 | cut -c1-170
done
for e in 0 75776 37888; do
HPPB200_WF_SLOTS=$e HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 2 > gpurun_out/wf_p.txt 2>&1
echo dich slots $e; grep -E "^n=|rounds=" gpurun_out/wf_p.txt | tail -2 | cut -c1-170
done
