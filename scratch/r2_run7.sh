cd $GRAFT_REPO_ROOT
python scratch/r2_dbg.py 2>&1 | grep -E "^reverse|edge" | cut -c1-60
python -m pytest tests -x -q -m gpu -k "wavefront or humanoid or freeflyer or full_size or birrt or dichotomy" > gpurun_out/r2_pytest_wf.txt 2>&1; tail -3 gpurun_out/r2_pytest_wf.txt
for c in 2 1; do
HPPB200_WF_CONTROL=$c HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 1 > gpurun_out/wf_c2.txt 2>&1
echo "control form $c"; grep -E "parity|^n=|rounds=" gpurun_out/wf_c2.txt | tail -4 | cut -c1-330
done
HPPB200_WF_TRACE=1 python tools/wf_run.py 1048576 20000 2 > gpurun_out/wf_d.txt 2>&1
echo "dichotomy"; grep -E "parity|^n=|rounds=" gpurun_out/wf_d.txt | tail -4 | cut -c1-330
HPPB200_WF_COMPACT=0 python tools/wf_run.py 1048576 0 2 > gpurun_out/wf_d0.txt 2>&1
echo "dichotomy no compaction"; grep -E "parity|^n=|rounds=" gpurun_out/wf_d0.txt | tail -1 | cut -c1-330
