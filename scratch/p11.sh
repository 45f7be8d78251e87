cd /root/repo
timeout 900 python bench.py > gpurun_out/r1t_bench.json 2> gpurun_out/r1t_bench.err; tail -n 2 gpurun_out/r1t_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1t_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['parity_vs_oracle_50k'], d['clocks']); print({k:(v['value'],v['e2e']) for k,v in d['edges'].items()})
for k,v in d['knn'].items():
    if isinstance(v,dict) and 'value' in v: print(k, round(v['value']), v.get('ms_per_call'))"
timeout 200 python scratch/tc2.py > gpurun_out/r1t_tc2_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"knnTcFilter|knnTcRerank" -s 5 -c 3 -o gpurun_out/r1t_knnTc python scratch/tc2.py > gpurun_out/r1t_ncu2.log 2>&1
timeout 200 python scratch/tc2.py > gpurun_out/r1t_tc2_plain2.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1t_knnTc_launches.csv python scratch/tc2.py > gpurun_out/r1t_ncu1.log 2>&1
ls gpurun_out | grep r1t
