cd $GRAFT_REPO_ROOT
for env in "" "HPPB200_TAIL=0" "HPPB200_TAIL_MAX=0"; do
echo "== $env"; env $env python tools/edge_rate.py humanoid_obst16.json 2 1048576 0.2 2>&1 | tail -1
done
for env in "" "HPPB200_ENGINE=0"; do
echo "== prog $env"; env $env python tools/edge_rate.py humanoid_obst16.json 1 1048576 0.2 2>&1 | tail -1
done
