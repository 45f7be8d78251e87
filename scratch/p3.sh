cd /root/repo
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3 > gpurun_out/r1l_pytest.log; cat gpurun_out/r1l_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1 | cut -c1-200
python bench.py > gpurun_out/r1l_bench.json 2> gpurun_out/r1l_bench.err; tail -n 3 gpurun_out/r1l_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1l_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value']); print({k:(v['value'],v['e2e']) for k,v in d['edges'].items()}); print(d['knn']['q1024_k10'], d['knn']['q1_k1'])"
python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4 2>&1 | grep method
HPPB200_WF_MIN=1000000000 python tools/edge_rate.py humanoid_obst6.json 2 1048576 0.4 2>&1 | grep method
