cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4 > gpurun_out/r1r_pytest.log; cat gpurun_out/r1r_pytest.log
timeout 900 python bench.py > gpurun_out/r1r_bench.json 2> gpurun_out/r1r_bench.err; tail -n 3 gpurun_out/r1r_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1r_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['parity_vs_oracle_50k']); print({k:(v['value'],v['e2e']) for k,v in d['edges'].items()})"
timeout 300 python bench.py --no-edges --no-cpu --steps 3 --warmup 3 > gpurun_out/r1r_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"configPassKernel|gjkJobsKernel" -s 6 -c 2 -o gpurun_out/r1r_config python bench.py --no-edges --no-cpu --steps 3 --warmup 3 > gpurun_out/r1r_ncu.log 2>&1
ls gpurun_out | grep r1r
