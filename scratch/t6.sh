cd /root/repo
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r1k_bench_n2.json 2> gpurun_out/r1k_bench_n2.err; python -c "
import json; d=json.loads(open('gpurun_out/r1k_bench_n2.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['edges_sharded'], d['edges']['progressive']['value'])"; head -c 300 gpurun_out/r1k_bench_n2.json; tail -n 3 gpurun_out/r1k_bench_n2.err
