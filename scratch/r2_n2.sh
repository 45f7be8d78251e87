cd $GRAFT_REPO_ROOT
nvidia-smi -L
python -m pytest tests -x -q -m gpu -k "multi_gpu or exchange" 2>&1 | tail -3
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --full-size > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err ) 2> gpurun_out/r2_bench_n2.time
grep -c "NCCL INFO" gpurun_out/r2_bench_n2.err; grep -E "Error|error|Traceback" gpurun_out/r2_bench_n2.err | head -5; tail -3 gpurun_out/r2_bench_n2.err | cut -c1-300; cat gpurun_out/r2_bench_n2.time
