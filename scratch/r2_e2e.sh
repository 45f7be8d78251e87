cd $GRAFT_REPO_ROOT
for pr in 1 0; do
echo "jobs stream high priority: $pr"; HPPB200_CFG_JOBS_PRIO=$pr python scratch/r2_e2e.py 1 2>&1 | grep -E "slices|False"
HPPB200_CFG_JOBS_PRIO=$pr python scratch/r2_e2e.py 1 2>&1 | grep -E "slices|False"
done
HPPB200_CFG_TRACE=1 python scratch/r2_e2e.py 1 2>&1 | grep -E "cfg\]" | head -27 | tail -7
