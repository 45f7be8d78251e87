cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4 > gpurun_out/r1q_pytest.log; cat gpurun_out/r1q_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1 | cut -c1-200
timeout 900 python bench.py > gpurun_out/r1q_bench.json 2> gpurun_out/r1q_bench.err; tail -n 3 gpurun_out/r1q_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1q_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac']); print(json.dumps(d['knn']['q1024_k10'])); print(json.dumps(d['knn']['q8192_k10']))"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 2>/dev/null | tail -n 1 | cut -c1-600
