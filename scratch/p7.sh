cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4 > gpurun_out/r1s_pytest.log; cat gpurun_out/r1s_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1 | cut -c1-160
timeout 900 python bench.py > gpurun_out/r1s_bench.json 2> gpurun_out/r1s_bench.err; tail -n 3 gpurun_out/r1s_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1s_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['traffic'], d['parity_vs_oracle_50k']); print({k:(v['value'],v['e2e']) for k,v in d['edges'].items()})
for k,v in d['knn'].items():
    if isinstance(v,dict) and 'value' in v: print(k, round(v['value']), v.get('ms_per_call'))"
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1s_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1s_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/r1s_ncu1.log 2>&1
ls gpurun_out | grep r1s
