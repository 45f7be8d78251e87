cd $GRAFT_REPO_ROOT
python -m pytest tests -x -q -m gpu -k "wavefront or humanoid or freeflyer or full_size or birrt or dichotomy" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -3 gpurun_out/r2_pytest_wf.txt
for l in 1 0; do
HPPB200_WF_LPT=$l HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_lpt$l.txt 2>&1
echo "prog lpt $l"; grep -E "parity|^n=|rounds=" gpurun_out/wf_lpt$l.txt | tail -4 | cut -c1-330
HPPB200_WF_LPT=$l HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 2 > gpurun_out/wf_d_lpt$l.txt 2>&1
echo "dichotomy lpt $l"; grep -E "parity|^n=|rounds=" gpurun_out/wf_d_lpt$l.txt | tail -4 | cut -c1-330
done
