cd $GRAFT_REPO_ROOT
for r in 32 24 16 8; do
HPPB200_WF_REFILL=$r HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_p.txt 2>&1
echo refill $r; grep -E "^n=|rounds=" gpurun_out/wf_p.txt | tail -2 | cut -c1-200
done
for a in 1.0 1.2 1.5 2.0; do
HPPB200_ENGINE_ALPHA=$a HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 0 1 > gpurun_out/wf_p.txt 2>&1
echo alpha $a; grep -E "^n=|rounds=" gpurun_out/wf_p.txt | tail -2 | cut -c1-330
done
