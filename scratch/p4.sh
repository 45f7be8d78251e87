cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4 > gpurun_out/r1p_pytest.log; cat gpurun_out/r1p_pytest.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 1 | cut -c1-200
timeout 900 python bench.py > gpurun_out/r1p_bench.json 2> gpurun_out/r1p_bench.err; tail -n 3 gpurun_out/r1p_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1p_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac']); print({k:(v['value'],v['e2e']) for k,v in d['edges'].items()}); print(json.dumps(d['knn'], indent=0))"
timeout 200 python scratch/tc2.py > gpurun_out/r1p_tc2_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r1p_knnTc_launches.csv python scratch/tc2.py > gpurun_out/r1p_ncu1.log 2>&1
timeout 200 python scratch/tc2.py > gpurun_out/r1p_tc2_plain2.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"knnTcFilter|knnTcRerank" -s 5 -c 3 -o gpurun_out/r1p_knnTc python scratch/tc2.py > gpurun_out/r1p_ncu2.log 2>&1
ls gpurun_out | grep r1p
