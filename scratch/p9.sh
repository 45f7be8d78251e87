cd /root/repo
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r1s_bench_n2.json 2> gpurun_out/r1s_bench_n2.err; echo "exit $?"; tail -n 3 gpurun_out/r1s_bench_n2.err | cut -c1-300; python -c "
import json; d=json.loads(open('gpurun_out/r1s_bench_n2.json').read().strip().splitlines()[-1]); print(d['n_gpus'], d['value'], d['e2e']['value'], d.get('edges_sharded')); print({k:v for k,v in (d.get('knn') or {}).items() if k.startswith('q1024')})" 2>&1 | cut -c1-1200
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 2>/dev/null | tail -n 1 | cut -c1-300
