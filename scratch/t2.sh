cd /root/repo
python -m pytest tests -m gpu -x -q -k wavefront 2>&1 | tail -n 15
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --no-edges > gpurun_out/r1h_bench_n2.json 2> gpurun_out/r1h_bench_n2.err; tail -c 1500 gpurun_out/r1h_bench_n2.json; tail -n 5 gpurun_out/r1h_bench_n2.err
