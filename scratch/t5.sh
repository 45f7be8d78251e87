cd /root/repo
python -m pytest tests -m gpu -x -q -k "knn" 2>&1 | tail -n 5
python bench.py --steps 5 --warmup 3 > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err; python -c "
import json; d=json.loads(open('gpurun_out/r1i_bench.json').read().strip().splitlines()[-1]); print(json.dumps(d['knn'], indent=1)); print(d['edges']['progressive'])"; tail -n 5 gpurun_out/r1i_bench.err
