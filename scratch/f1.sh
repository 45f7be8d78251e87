cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 > gpurun_out/r1m_pytest.log; cat gpurun_out/r1m_pytest.log
cat > /tmp/knnprof.py <<'PY'
import sys; sys.path.insert(0,'/root/repo')
import numpy as np, __graft_entry__ as ge
pkg=ge.load(); desc=pkg.ModelDescription.load("ur5_boxes8.json"); m=pkg.Model(desc)
nn=pkg.NearestNeighbor(m,0); nn.append(pkg.uniform_configs(desc,1<<20,seed=3))
qs=pkg.uniform_configs(desc,1024,seed=4)
for _ in range(2): nn.query(qs,10)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:knnTileTopk --launch-skip 1 --launch-count 1 -o gpurun_out/r1m_knn -f python /tmp/knnprof.py > gpurun_out/r1m_knn.log 2>&1
ls -la gpurun_out/r1m_knn*
