cd /root/repo
timeout 900 python bench.py > gpurun_out/r1u_bench.json 2> gpurun_out/r1u_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1u_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['parity_vs_oracle_50k'])
for k,v in d['knn'].items():
    if isinstance(v,dict) and 'value' in v: print(k, round(v['value']), v.get('ms_per_call'))"
